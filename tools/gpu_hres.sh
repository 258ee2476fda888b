#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 -k "conv" 2>&1 | tail -5
for v in 0 1; do echo "NO_HRES=$v"; PL_IGEMM_NO_HRES=$v timeout 120 python tools/layer_bench.py r2_3x3_64 r3_3x3_128 r4_3x3_256 2>&1 | grep b256; done
