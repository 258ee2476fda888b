#!/bin/bash
# sub-batched schedules (DeviceNet.bind schedule / PL_SUBBATCH) on ResNet-50 b256: parity first, then timing
mkdir -p gpurun_out
PL_SUBBATCH=0-12:2,13-30:2 timeout 600 python -m pytest tests/test_gpu_nets.py -q -x 2>&1 | tail -3
for sb in "" "0-12:4" "0-12:8" "3-12:4" "0-25:4" "0-12:4,13-25:2" "0-2:8,3-12:4" "0-2:4" "0-12:2"; do
  PL_SUBBATCH=$sb timeout 600 python bench.py --net resnet50 --steps 30 --warmup 5 --no-cpu-baseline 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PL_SUBBATCH=$sb', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'], 'launches', l['gpu_launches']//l['steps'])"
done | tee gpurun_out/subbatch.txt
