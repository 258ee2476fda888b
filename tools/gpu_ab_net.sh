#!/bin/bash
# same-box A/B of whole nets: HEAD build (variants/libpl_base.so) against the working tree's library
BASE=${PL_AB_BASE:-$PWD/paddle-lite_b200/variants/libpl_base.so}
run() { python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d.get('config',{}).get('workload'), d['ms_per_step'], 'ms', round(d['value']), d['unit'])"; }
for i in 1 2; do
  echo "--- base"; PL_INT8_LIB=$BASE run --net resnet50 --batch 256; PL_INT8_LIB=$BASE run --net mobilenet_v1 --batch 64
  echo "--- new"; run --net resnet50 --batch 256; run --net mobilenet_v1 --batch 64
done
