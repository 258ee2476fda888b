#!/bin/bash
# what bounds the tile period of the short-K layers?  debug build (PL_DEBUG_KNOBS): drop one phase at a time
# (1 epilogue math, 2 MMAs, 8 halo loads, 16 TMA stores) and compare the layer time at batch 256.
export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so
for l in r2_3x3_64 r2_1x1_64_256 r2_1x1_64_256_res r3_3x3_128 stem7x7_s2d r2_1x1_64_64; do
  for f in 0 16 8 24 1 2 17; do
    echo -n "flags=$f "
    PL_DBG_FLAGS=$f timeout 120 python tools/layer_bench.py $l 2>&1 | tail -1 | cut -c1-110
  done
done
