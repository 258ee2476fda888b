#!/bin/bash
# 128-wide tiles (4 accumulator stages) with two epilogue groups on alternate tiles
L="r2_1x1_64_256 r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res"
echo "--- BN 256, one group"; PL_EPI_G2=0 python tools/layer_bench.py $L 2>&1 | cut -c1-120
echo "--- BN 128, two groups"; PL_FORCE_BN=128 PL_IGEMM_RES_BN=128 PL_EPI_G2_MIN_UNITS=8 python tools/layer_bench.py $L 2>&1 | cut -c1-120
echo "--- BN 128, two groups, 6 res bufs"; PL_FORCE_BN=128 PL_IGEMM_RES_BN=128 PL_EPI_G2_MIN_UNITS=8 PL_IGEMM_RES_BUFS=6 PL_IGEMM_OUT_BUFS=4 python tools/layer_bench.py $L 2>&1 | cut -c1-120
M="r3_3x3_128 r3_1x1_256_128 r3_1x1_512_128 r5_1x1_512_2048_res m_pw_14_512_512"
echo "--- 128-wide layers, one group"; PL_EPI_G2=0 python tools/layer_bench.py $M 2>&1 | cut -c1-120
echo "--- 128-wide layers, two groups"; PL_EPI_G2_MIN_UNITS=8 PL_EPI_G2_MAX_K=100000 python tools/layer_bench.py $M 2>&1 | cut -c1-120
