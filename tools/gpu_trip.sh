#!/bin/bash
# One GPU trip: diagnostics first, then the parity suite split into processes so
# that one trapped kernel does not hide the state of everything else.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python tools/diag.py > gpurun_out/diag.log 2>&1; echo "diag rc=$?" >> gpurun_out/diag.log
for grp in "gemm" "conv_int32 or kats" "conv_int8 or conv_fp32 or legacy or full_size or slice or error" \
           "depthwise" "calib or rounding or layout or pool or add or peak"; do
  name=$(echo "$grp" | tr ' ' '_')
  timeout 900 python -m pytest tests -m gpu -q --timeout 300 -k "$grp" > "gpurun_out/pytest_${name}.log" 2>&1
  echo "rc=$?" >> "gpurun_out/pytest_${name}.log"
done
tail -n 30 gpurun_out/diag.log
for f in gpurun_out/pytest_*.log; do echo "== $f"; tail -n 8 "$f"; done
