#!/bin/bash
# conv/gemm parity then the main layer timings
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 -k "conv or gemm or thin" 2>&1 | tail -3
timeout 300 python tools/layer_bench.py ${@:-r2_3x3_64 r3_3x3_128 r4_3x3_256 r5_3x3_512 r2_1x1_64_256_res r3_1x1_128_512_res r4_1x1_256_1024_res r5_1x1_512_2048_res r2_1x1_256_64 r2_1x1_64_64 r4_1x1_1024_256 stem7x7} 2>&1 | grep b256
