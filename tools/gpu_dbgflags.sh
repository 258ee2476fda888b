#!/bin/bash
for f in 0 1 2 3 4 6 7 8 12 14 15; do echo "flags=$f"; PL_DBG_FLAGS=$f timeout 120 python tools/layer_bench.py r2_3x3_64 r3_3x3_128 2>&1 | grep b256; done
