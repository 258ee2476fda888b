#!/bin/bash
# ablation of one layer with the debug-knob build: which role bounds the tile period
# flags: 1 skip epilogue math, 2 skip MMAs, 16 skip output stores, 4/8 producer variants
export PL_INT8_LIB=paddle-lite_b200/variants/libpl_dbg.so
for L in "$@"; do
  for F in ${FLAGS:-0 16 1 2 17 19}; do
    echo -n "flags=$F "; PL_DBG_FLAGS=$F python tools/layer_bench.py $L --iters 20 --nsets 1 $EXTRA 2>&1 | grep "ms="
  done
done
