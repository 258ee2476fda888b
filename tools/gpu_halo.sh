#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 2>&1 | tail -4
PL_IGEMM_NO_HALO=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 300 -k "conv" 2>&1 | tail -1
timeout 600 python -m pytest tests/test_gpu_nets.py tests/test_gpu_plugin.py -m gpu -q -x --timeout 300 2>&1 | tail -2
python tools/layer_bench.py r2_3x3_64 r3_3x3_128 r4_3x3_256 r5_3x3_512 | grep b256
