#!/bin/bash
L="r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res r5_1x1_512_2048_res r2_1x1_256_64 r2_3x3_64 r2_1x1_64_256"
run() { echo "== $*"; env "$@" timeout 300 python tools/layer_bench.py $L 2>&1 | grep "b256" ; }
run A=default
run PL_IGEMM_OUT_BUFS=1
run PL_IGEMM_RES_BN=128 PL_IGEMM_RES_BK=128
run PL_IGEMM_RES_BN=256 PL_IGEMM_RES_BK=64
run PL_IGEMM_RES_BN=256 PL_IGEMM_RES_BK=128
run PL_IGEMM_RES_BUFS=2
run PL_IGEMM_RES_BUFS=4
run PL_IGEMM_MIN_STAGES=2
run PL_IGEMM_MIN_STAGES=4
