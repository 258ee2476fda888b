#!/bin/bash
# full GPU suite, then the round profile of the resnet50 bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_all.log
tail -n 8 gpurun_out/pytest_all.log
bash tools/gpu_profile_round.sh resnet50 r01
