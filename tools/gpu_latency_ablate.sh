#!/bin/bash
# where do the ~5.5 us of a one-tile conv launch go?  debug build (PL_DEBUG_KNOBS): skip phases and compare.
export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so
for f in 0 31 4096 5120 4608 5632 8192; do
  echo -n "PL_DBG_FLAGS=$f: "
  PL_DBG_FLAGS=$f timeout 120 python tools/conv1_latency.py 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(d['us_per_launch'])"
done
