#!/bin/bash
# ONE ncu --set full invocation over the timed launches of a few layers at BASELINE shapes.
# usage: gpu_ncu_full.sh tag layer [layer...]
TAG=$1; shift
mkdir -p gpurun_out
python tools/layer_bench.py "$@" --iters 2 --nsets 1 > gpurun_out/${TAG}_plain_layers.log 2>&1 || { tail -5 gpurun_out/${TAG}_plain_layers.log; exit 1; }
cat gpurun_out/${TAG}_plain_layers.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"igemm_i8_kernel|dw3x3|pool4" -c 60 -f -o gpurun_out/${TAG}_full \
   python tools/layer_bench.py "$@" --iters 2 --nsets 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/${TAG}_ncu_full.log; ls -la gpurun_out/${TAG}_full.ncu-rep
