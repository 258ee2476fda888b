#!/bin/bash
# same-box A/B of the working tree's library against a variant (default: variants/libpl_base.so = HEAD build)
# usage: gpu_ab_layers.sh [layer ...]
BASE=${PL_AB_BASE:-$PWD/paddle-lite_b200/variants/libpl_base.so}
L="$@"
[ -z "$L" ] && L="r2_3x3_64 r3_3x3_128 r4_3x3_256 r5_3x3_512 stem7x7_s2d stem3x3_s2d r2_1x1_64_64 r2_1x1_64_256 r2_1x1_64_256_res"
echo "--- base"; PL_INT8_LIB=$BASE python tools/layer_bench.py $L 2>&1 | cut -c1-100
echo "--- new";  python tools/layer_bench.py $L 2>&1 | cut -c1-100
echo "--- base"; PL_INT8_LIB=$BASE python tools/layer_bench.py $L 2>&1 | cut -c1-100
echo "--- new";  python tools/layer_bench.py $L 2>&1 | cut -c1-100
