#!/bin/bash
# 8-GPU sanity: batch-sharded bench, reference arm under torchrun, multi-GPU tests, N-split GEMM sweep
mkdir -p gpurun_out/r01_8gpu
O=gpurun_out/r01_8gpu
N=${1:-8}
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $O/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 > $O/bench_resnet50.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 1500 $O/bench_resnet50.json
timeout 600 $TR bench.py --gpus $N --net mobilenet_v1 --steps 30 --warmup 5 --no-cpu-baseline > $O/bench_mobilenet.json 2>> $O/bench.err; echo "bench mb rc=$?"
timeout 600 $TR tools/gemm_sweep.py 4096x4096x4096 8192x8192x8192 16384x16384x16384 64x16384x16384 > $O/gemm_sweep.jsonl 2> $O/gemm.err; echo "gemm rc=$?"; cat $O/gemm_sweep.jsonl
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 500 2>&1 | tail -3
