#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_plugin.py tests/test_gpu_nets.py tests/test_gpu_cta2.py -q -x 2>&1 | tail -2
timeout 300 python tools/mma_probe.py 2>&1 | grep -E "pairs|row_bytes=128"
L="r5_1x1_512_2048_res m_pw_14_512_512 r3_1x1_512_128 r4_1x1_256_1024_res r5_1x1_2048_512"
for v in 1 0; do echo "== PL_NO_BN128_RES=$v"; PL_NO_BN128_RES=$v timeout 300 python tools/layer_bench.py $L --iters 50 2>&1 | grep b256; done
for v in 1 0; do echo "== PL_NO_BN128_RES=$v b64"; PL_NO_BN128_RES=$v timeout 300 python tools/layer_bench.py m_pw_14_512_512 --batch 64 --iters 50 2>&1 | grep b64; done
echo "== fc shapes (gemv up to M=64)"
for v in 1 0; do echo "-- PL_NO_GEMV=$v";  PL_NO_GEMV=$v timeout 300 python tools/gemm_sweep.py 64x1000x1024 32x1000x1024 16x4096x4096 64x4096x4096 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['shape'], d['out'], d['ms_gemm'], d['tops_gemm'])"; done
for net in mobilenet_v1 resnet50; do
  timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
done
