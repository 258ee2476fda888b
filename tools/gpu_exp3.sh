#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_cta2.py tests/test_gpu_parity.py -q -x 2>&1 | tail -2
PL_CTA2_HRES=2 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_nets.py -q -x 2>&1 | tail -2
L="r2_3x3_64 r3_3x3_128 r4_3x3_256"
for v in 0 1 2; do echo "== PL_CTA2_HRES=$v"; PL_CTA2_HRES=$v timeout 300 python tools/layer_bench.py $L --iters 50 2>&1 | grep b256; done
for v in 0 1 2; do
  PL_CTA2_HRES=$v timeout 600 python bench.py --net resnet50 --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PL_CTA2_HRES=$v resnet50', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
done
