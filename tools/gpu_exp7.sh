#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for v in 1 0; do
  echo "== PL_KGROUP=$v (1 = one k-block per barrier, 0 = auto)"
  PL_KGROUP=$v timeout 300 python tools/layer_bench.py r2_1x1_64_256_res r2_1x1_256_64 r3_1x1_512_128 r3_1x1_256_128 r3_1x1_128_512_res r4_1x1_1024_256 r4_1x1_256_1024_res r5_1x1_2048_512 r5_3x3_512 r3_3x3s2_128 r4_3x3s2_256 m_pw_14_512_512 --iters 50 2>&1 | grep b256 | cut -c1-160
  PL_KGROUP=$v timeout 300 python tools/layer_bench.py m_pw_14_512_512 r3_1x1_512_128 --batch 64 --iters 50 2>&1 | grep " b64 " | cut -c1-160
  PL_KGROUP=$v timeout 300 python tools/gemm_sweep.py 2048x2048x2048 4096x4096x4096 8192x8192x8192 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['shape'], d['out'], d['ms_gemm'], d['tops_gemm'], d['frac_mma_peak'])"
  for net in mobilenet_v1 resnet50 humanseg; do
  PL_KGROUP=$v timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
  done
done
