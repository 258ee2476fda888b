#!/bin/bash
# A/B of the 256x256 tile-pair GEMM mode: parity tests, then the large-GEMM sweep with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -x -k "gemm" 2>&1 | tail -4
S="2048x2048x2048 4096x4096x4096 8192x8192x8192 16384x16384x16384 8192x8192x2048 4096x16384x4096"
echo "== pair mode"; timeout 600 python tools/gemm_sweep.py $S 2>&1 | grep '"int8"' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print({k:d[k] for k in d if k in ('shape','ms_gemm','tops_gemm','frac_mma_peak')})"
echo "== staged pair"; PL_MPAIR_STAGED=1 timeout 600 python tools/gemm_sweep.py $S 2>&1 | grep '"int8"' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print({k:d[k] for k in d if k in ('shape','ms_gemm','tops_gemm','frac_mma_peak')})"
