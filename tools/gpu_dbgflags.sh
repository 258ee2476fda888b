#!/bin/bash
# usage: gpu_dbgflags.sh "layers" flags...
L=$1; shift
for f in "$@"; do echo "flags=$f"; PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so PL_DBG_FLAGS=$f timeout 120 python tools/layer_bench.py $L 2>&1 | grep b256; done
