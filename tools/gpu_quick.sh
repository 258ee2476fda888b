#!/bin/bash
# quick check of a change: GPU suite, both network benches (short), optional layer list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -x 2>&1 | tail -4
for net in resnet50 mobilenet_v1; do
  timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --breakdown gpurun_out/bd_$net.json 2> gpurun_out/q_err_$net.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'], 'e2e %.0f'%l['e2e']['value'], l['clocks']['sm_mhz'], l['roofline']['by_kind_ms'])"
  tail -3 gpurun_out/q_err_$net.log
done
if [ $# -gt 0 ]; then python tools/layer_bench.py "$@" --iters 20; fi
