#!/bin/bash
# Round profile, one gpurun call: the plain bench (exit 0 first), then ONE ncu pass over the same
# command with per-launch duration + DRAM bytes.  tools/make_traffic.py turns the launch list into
# profiles/traffic.json + the markdown summary.   usage: gpu_profile_round.sh [net] [tag]
NET=${1:-resnet50}
TAG=${2:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt
timeout 900 python bench.py --net $NET --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain_$NET.log 2>&1 || { tail -5 gpurun_out/${TAG}_plain_$NET.log; exit 1; }
tail -c 600 gpurun_out/${TAG}_plain_$NET.log
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -c 4000 --csv --log-file gpurun_out/${TAG}_launches_$NET.csv \
    python bench.py --net $NET --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_$NET.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/${TAG}_launches_$NET.csv
