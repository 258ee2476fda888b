#!/bin/bash
# MobileNetV1 b64 (and b256) A/B of the working tree's library against variants/libpl_base.so
BASE=$PWD/paddle-lite_b200/variants/libpl_base.so
run() { python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-extras --net mobilenet_v1 "$@" 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'): d=json.loads(l); print(d['ms_per_step'], round(d['value']))"; }
for i in 1 2; do
  echo -n "base b64: "; PL_INT8_LIB=$BASE run --batch 64
  echo -n "new  b64: "; run --batch 64
done
echo -n "base b256: "; PL_INT8_LIB=$BASE run --batch 256
echo -n "new  b256: "; run --batch 256
