#!/bin/bash
# epilogue-bound expand layers: 128-wide tiles (4 accumulator stages) against the default 256 (2 stages)
L="r2_1x1_64_256 r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res r5_1x1_512_2048_res"
echo "--- default"; python tools/layer_bench.py $L 2>&1 | cut -c1-200
echo "--- BN 128"; PL_FORCE_BN=128 PL_IGEMM_RES_BN=128 python tools/layer_bench.py $L 2>&1 | cut -c1-200
echo "--- BN 128, res bufs 6"; PL_FORCE_BN=128 PL_IGEMM_RES_BN=128 PL_IGEMM_RES_BUFS=6 PL_IGEMM_OUT_BUFS=4 python tools/layer_bench.py $L 2>&1 | cut -c1-200
echo "--- BN 64"; PL_FORCE_BN=64 PL_IGEMM_RES_BN=64 python tools/layer_bench.py r2_1x1_64_256 r2_1x1_64_256_res 2>&1 | cut -c1-200
