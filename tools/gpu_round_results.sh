#!/bin/bash
# Round results on one B200: every BASELINE config that fits one GPU.  usage: gpu_round_results.sh TAG
TAG=${1:-r01}
O=gpurun_out/${TAG}_results
mkdir -p $O
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,power.limit --format=csv > $O/gpu.txt
timeout 300 python tools/conv1_latency.py > $O/config1_conv.jsonl 2> $O/config1.err
timeout 900 python tools/layer_bench.py --iters 50 > $O/layers_b256.txt 2> $O/layers.err
timeout 600 python bench.py --net mobilenet_v1 --batch 1 --steps 200 --warmup 20 --no-cpu-baseline > $O/bench_mobilenet_b1.json 2> $O/m1.err
timeout 600 python bench.py --net mobilenet_v1 --steps 100 --warmup 10 --breakdown $O/bd_mobilenet_b64.json > $O/bench_mobilenet_b64.json 2> $O/m64.err
timeout 600 python bench.py --net shufflenet_v2_like --steps 100 --warmup 10 --no-cpu-baseline > $O/bench_shufflenet_b64.json 2> $O/s64.err
timeout 300 python tools/channel_ops_bench.py > $O/channel_ops.jsonl 2> $O/chan.err
timeout 600 python bench.py --steps 30 --warmup 5 --breakdown $O/bd_resnet50_b256.json > $O/bench_resnet50_b256.json 2> $O/r256.err
timeout 900 python tools/gemm_sweep.py > $O/gemm_sweep_1gpu.jsonl 2> $O/gemm.err
timeout 300 python tools/mma_probe.py > $O/mma_probe.txt 2>&1
tail -n 3 $O/*.err | tail -n 30
cat $O/config1_conv.jsonl
