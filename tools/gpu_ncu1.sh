#!/bin/bash
mkdir -p gpurun_out
L="r2_1x1_64_256 r2_1x1_64_256_res r2_3x3_64 stem7x7"
python tools/layer_bench.py > gpurun_out/layers.log 2>&1; cat gpurun_out/layers.log
python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:igemm -s 12 -c 8 -o gpurun_out/prof_igemm_r1 \
   python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/ncu.log 2>&1
tail -5 gpurun_out/ncu.log
