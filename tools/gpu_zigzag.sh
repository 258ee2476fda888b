#!/bin/bash
# alternate tile order (last-to-first on every second igemm launch) against the forward order, whole nets
run() { python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], 'ms', round(d['value']), d['unit'], 'traffic/launch', d.get('roofline',{}).get('traffic'))"; }
for i in 1 2; do
  echo "--- forward"; run --net resnet50 --batch 256; run --net mobilenet_v1 --batch 64
  echo "--- zigzag"; PL_TILE_ZIGZAG=1 run --net resnet50 --batch 256; PL_TILE_ZIGZAG=1 run --net mobilenet_v1 --batch 64
done
