#!/bin/bash
# ONE ncu --set full invocation over the timed launches of a few layers at BASELINE shapes; the summaries are
# made on the box (gpurun brings back at most 64 MiB: the .ncu-rep itself is dropped when it is larger than 45 MB).
# usage: gpu_ncu_full.sh tag layer [layer...]
TAG=$1; shift
mkdir -p gpurun_out
python tools/layer_bench.py "$@" --iters 2 --nsets 1 > gpurun_out/${TAG}_plain_layers.log 2>&1 || { tail -5 gpurun_out/${TAG}_plain_layers.log; exit 1; }
cat gpurun_out/${TAG}_plain_layers.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"igemm_i8_kernel|dw3x3|pool4" -c 60 -f -o gpurun_out/${TAG}_full \
   python tools/layer_bench.py "$@" --iters 2 --nsets 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/${TAG}_ncu_full.log; ls -la gpurun_out/${TAG}_full.ncu-rep
python tools/ncu_brief.py gpurun_out/${TAG}_full.ncu-rep > gpurun_out/${TAG}_brief.md 2> gpurun_out/${TAG}_brief.err
for i in 2 7 12 17 22 27 32 37; do python tools/ncu_hot.py gpurun_out/${TAG}_full.ncu-rep $i 14 >> gpurun_out/${TAG}_hot.txt 2>/dev/null; echo >> gpurun_out/${TAG}_hot.txt; done
SZ=$(stat -c %s gpurun_out/${TAG}_full.ncu-rep)
if [ "$SZ" -gt 45000000 ]; then rm -f gpurun_out/${TAG}_full.ncu-rep; echo "report dropped ($SZ bytes), summaries kept"; fi
