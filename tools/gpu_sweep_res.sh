#!/bin/bash
L="r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res r5_1x1_512_2048_res"
for bn in 128 256; do for nb in 2 3 4 6; do
  echo "== RES_BN=$bn RES_BUFS=$nb"
  PL_IGEMM_RES_BN=$bn PL_IGEMM_RES_BUFS=$nb timeout 300 python tools/layer_bench.py $L 2>&1 | grep -v "^$" | tail -4
done; done
echo "== old direct path"; PL_IGEMM_NO_TMA_EPI=1 timeout 300 python tools/layer_bench.py $L 2>&1 | tail -4
