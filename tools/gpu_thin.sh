#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "thin" --timeout 300 2>&1 | tail -8
timeout 600 python -m pytest tests/test_gpu_nets.py tests/test_gpu_plugin.py -m gpu -q -x --timeout 300 2>&1 | tail -3
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/breakdown_resnet50.json 2>&1 | tail -1 | cut -c1-160
timeout 600 python bench.py --net mobilenet_v1 --steps 10 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/breakdown_mobilenet.json 2>&1 | tail -1 | cut -c1-160
