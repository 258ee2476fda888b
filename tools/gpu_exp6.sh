#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_nets.py tests/test_gpu_model.py tests/test_gpu_cta2.py -q -x 2>&1 | tail -2
for v in 0 1; do
  echo "== PL_HALO_KEEP_NTILE=$v"
  PL_HALO_KEEP_NTILE=$v timeout 120 python tools/conv1_latency.py 2>/dev/null | cut -c1-110
  for b in 1 8 32; do PL_HALO_KEEP_NTILE=$v timeout 120 python tools/layer_bench.py r2_3x3_64 r3_3x3_128 r4_3x3_256 --batch $b --iters 50 2>&1 | grep " b$b " | cut -c1-150; done
  for net in humanseg mobilenet_v1 resnet50; do
  PL_HALO_KEEP_NTILE=$v timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
  done
done
