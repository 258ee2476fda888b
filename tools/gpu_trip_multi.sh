#!/bin/bash
# 2-GPU trip: NCCL N-split test, N-split sweep, batch-sharded bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x 2>&1 | tail -15
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/gemm_sweep.py 4096x4096x4096 8192x8192x8192 16384x16384x16384 64x16384x16384 > gpurun_out/gemm_sweep_2gpu.jsonl 2> gpurun_out/gemm_sweep_2gpu.err
tail -3 gpurun_out/gemm_sweep_2gpu.err; cat gpurun_out/gemm_sweep_2gpu.jsonl
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
tail -3 gpurun_out/bench_2gpu.err; cat gpurun_out/bench_2gpu.json
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_1gpu_same_box.json 2>/dev/null; cat gpurun_out/bench_1gpu_same_box.json
