#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_plugin.py tests/test_gpu_nets.py -m gpu -q -x -k "depthwise or dw or mobilenet or Mobile" 2>&1 | tail -5
timeout 300 python tools/layer_bench.py dw_112_32_s1 dw_112_64_s2 dw_56_128_s1 dw_14_512_s1 --batch 64 2>&1 | grep b64
timeout 600 python bench.py --net mobilenet_v1 --steps 10 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/breakdown_mobilenet.json 2>&1 | tail -1 | cut -c1-200
