#!/bin/bash
# fixed vs per-tile cost: same layer at several batch sizes.  usage: gpu_batchscale.sh "layers" "flags" batches...
L=$1; F=$2; shift; shift
for f in $F; do for b in "$@"; do echo -n "flags=$f "; PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so PL_DBG_FLAGS=$f timeout 120 python tools/layer_bench.py $L --batch $b 2>&1 | grep " b$b "; done; done
