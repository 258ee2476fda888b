#!/bin/bash
# L2 eviction hints (PL_L2_HINTS: bit 0 residual loads evict_first, bit 1 A rows of single-N-tile launches) on whole nets
run() { python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], 'ms', round(d['value']), d['unit'])"; }
for h in 0 15 31 47 63 15 0; do
  echo "--- PL_L2_HINTS=$h"; PL_L2_HINTS=$h run --net resnet50 --batch 256
done
for h in 0 15 31 0 15 31; do echo "mobilenet PL_L2_HINTS=$h"; PL_L2_HINTS=$h run --net mobilenet_v1 --batch 64; done
