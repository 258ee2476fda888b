#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
L="stem7x7_s2d stem3x3_s2d r2_3x3_64 r2_1x1_64_64 r2_1x1_256_64 m_pw_112_32_64 r2_1x1_64_256_res"
timeout 300 python tools/layer_bench.py $L --iters 50 2>&1 | grep b256 | cut -c1-150
for net in mobilenet_v1 resnet50 humanseg; do
  timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
done
