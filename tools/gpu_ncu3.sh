#!/bin/bash
# usage: gpu_ncu3.sh layer [layer...]   (ncu --set full of the timed launch of each layer)
mkdir -p gpurun_out
for L in "$@"; do
python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/plain_$L.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"igemm|dw|pool|im2col" -s 3 -c 1 -f -o gpurun_out/prof_$L \
   python tools/layer_bench.py $L --iters 2 --nsets 1 > gpurun_out/ncu_$L.log 2>&1
tail -2 gpurun_out/ncu_$L.log
done
