#!/bin/bash
# usage: gpu_trace.sh layer [flags]   -> gpurun_out/trace_<layer>_<flags>.txt
mkdir -p gpurun_out
L=$1; F=${2:-0}
PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so PL_DBG_FLAGS=$F PL_DBG_TRACE=gpurun_out/trace_${L}_$F.txt timeout 120 python tools/layer_bench.py $L --iters 2 --nsets 1 2>&1 | grep b256
