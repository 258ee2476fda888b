#!/bin/bash
# per-op tile order from the dependency chain (nets.DeviceNet.bind, PL_NET_ZIGZAG) against first-to-last everywhere
run() { python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras "$@" 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], 'ms', round(d['value']), d['unit'])"; }
for i in 1 2; do
  echo "--- forward"; PL_NET_ZIGZAG=0 run --net resnet50 --batch 256; PL_NET_ZIGZAG=0 run --net mobilenet_v1 --batch 64; PL_NET_ZIGZAG=0 run --net shufflenet_v2_like --batch 64
  echo "--- zigzag"; run --net resnet50 --batch 256; run --net mobilenet_v1 --batch 64; run --net shufflenet_v2_like --batch 64
done
