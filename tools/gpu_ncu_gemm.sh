#!/bin/bash
# ncu --set full of the large int8 GEMM (config 5): banded tile pairs vs row-major pairs vs single tiles.  usage: gpu_ncu_gemm.sh TAG
TAG=${1:-r01_gemm}
mkdir -p gpurun_out
for mode in banded rowmajor single; do
  unset PL_NO_MPAIR PL_GROUP_M
  if [ $mode = single ]; then export PL_NO_MPAIR=1 PL_GROUP_M=1; fi
  if [ $mode = rowmajor ]; then export PL_GROUP_M=1; fi
  timeout 300 python tools/gemm_sweep.py 8192x8192x8192 2>&1 | grep '"int8"' | cut -c1-160
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"igemm_i8_kernel" -s 8 -c 2 -f -o gpurun_out/${TAG}_$mode \
     python tools/gemm_sweep.py 8192x8192x8192 > gpurun_out/${TAG}_${mode}_ncu.log 2>&1
  echo "ncu $mode rc=$?"
  python tools/ncu_brief.py gpurun_out/${TAG}_$mode.ncu-rep > gpurun_out/${TAG}_${mode}_brief.md 2>/dev/null
  python tools/ncu_hot.py gpurun_out/${TAG}_$mode.ncu-rep 0 12 > gpurun_out/${TAG}_${mode}_hot.txt 2>/dev/null
  SZ=$(stat -c %s gpurun_out/${TAG}_$mode.ncu-rep); if [ "$SZ" -gt 20000000 ]; then rm -f gpurun_out/${TAG}_$mode.ncu-rep; fi
  cat gpurun_out/${TAG}_${mode}_brief.md
done
