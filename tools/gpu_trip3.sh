#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_all.log
tail -n 15 gpurun_out/pytest_all.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/breakdown_resnet50.json > gpurun_out/bench_resnet50.log 2>&1; echo "rc=$?" >> gpurun_out/bench_resnet50.log
tail -n 4 gpurun_out/bench_resnet50.log
timeout 600 python bench.py --net mobilenet_v1 --steps 10 --warmup 3 --no-cpu-baseline --breakdown gpurun_out/breakdown_mobilenet.json > gpurun_out/bench_mobilenet.log 2>&1; echo "rc=$?" >> gpurun_out/bench_mobilenet.log
tail -n 4 gpurun_out/bench_mobilenet.log
