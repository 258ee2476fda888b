#!/bin/bash
# CTA-0 role timelines of a few layers (needs the dbg variant: tools/build_variant.sh dbg "-DPL_DEBUG_KNOBS")
mkdir -p gpurun_out
export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so
for l in "$@"; do
  PL_DBG_TRACE_ON=1 PL_DBG_TRACE=gpurun_out/trace_$l.txt timeout 120 python tools/layer_bench.py $l --iters 1 --nsets 1 > /dev/null 2>&1
  echo "=== $l"; python tools/trace_timeline.py gpurun_out/trace_$l.txt 10 4
done
