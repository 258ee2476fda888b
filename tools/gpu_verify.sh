#!/bin/bash
# Verify HEAD on one B200: whole GPU suite, then the two network benches with per-layer breakdowns.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
( time timeout 1200 python -m pytest tests -m gpu -q --timeout 300 ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --breakdown gpurun_out/bd_resnet50.json > gpurun_out/bench_resnet50.json 2> gpurun_out/bench_resnet50.err; echo "bench rc=$?"
cat gpurun_out/bench_resnet50.json
timeout 600 python bench.py --net mobilenet_v1 --steps 50 --warmup 5 --no-cpu-baseline --breakdown gpurun_out/bd_mobilenet.json > gpurun_out/bench_mobilenet.json 2> gpurun_out/bench_mobilenet.err; echo "bench rc=$?"
cat gpurun_out/bench_mobilenet.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>&1; cat gpurun_out/bench_reference.json
