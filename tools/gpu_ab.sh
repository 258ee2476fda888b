#!/bin/bash
# A/B of an environment switch on both network benches:  gpu_ab.sh VAR valA valB [extra bench args]
VAR=$1; A=$2; B=$3; shift 3
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -x 2>&1 | tail -3
for v in $A $B $A $B; do
  for net in resnet50 mobilenet_v1; do
    env $VAR=$v timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline "$@" 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$VAR=$v', '$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'], 'e2e %.0f'%l['e2e']['value'], l['clocks']['sm_mhz'])"
  done
done
