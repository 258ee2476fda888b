#!/bin/bash
# two half-width epilogue groups on alternate tiles (PL_EPI_G2) on the wide short-K layers, A/B on one box
L="r2_1x1_64_256 r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res r2_1x1_64_64 stem7x7_s2d r3_1x1_256_128 m_pw_112_32_64"
echo "--- PL_EPI_G2=0"; PL_EPI_G2=0 python tools/layer_bench.py $L 2>&1 | cut -c1-120
echo "--- default (G2 for K <= 256, 16 units)"; python tools/layer_bench.py $L 2>&1 | cut -c1-120
echo "--- G2 also for 8 units"; PL_EPI_G2_MIN_UNITS=8 python tools/layer_bench.py $L 2>&1 | cut -c1-120
echo "--- G2 any K"; PL_EPI_G2_MAX_K=100000 python tools/layer_bench.py r3_1x1_256_128 r5_1x1_512_2048_res r4_1x1_1024_256 r3_1x1_512_128 2>&1 | cut -c1-120
echo "--- G2=0 ref"; PL_EPI_G2=0 python tools/layer_bench.py r3_1x1_256_128 r5_1x1_512_2048_res r4_1x1_1024_256 r3_1x1_512_128 2>&1 | cut -c1-120
