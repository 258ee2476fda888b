#!/bin/bash
mkdir -p gpurun_out
for L in r2_1x1_64_256 r2_1x1_64_256_res r2_1x1_256_64; do
python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/plain_$L.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:igemm -s 3 -c 1 -o gpurun_out/prof_$L \
   python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/ncu_$L.log 2>&1
tail -2 gpurun_out/ncu_$L.log
done
