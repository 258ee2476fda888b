#!/bin/bash
# Final evidence of a round: results + ncu launch lists (both nets) + the reference arm.  (ncu --set full: gpu_ncu_full.sh,
# a call of its own -- gpurun copies back at most 64 MiB.)
TAG=${1:-r01}
bash tools/gpu_round_results.sh $TAG
bash tools/gpu_profile_round.sh resnet50 $TAG
bash tools/gpu_profile_round.sh mobilenet_v1 $TAG
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>&1
tail -2 gpurun_out/${TAG}_bench_reference.json | cut -c1-400
du -sh gpurun_out
