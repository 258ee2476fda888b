#!/bin/bash
# sweep one environment knob over a list of layers:  gpu_sweep_env.sh VAR "v1 v2 ..." layer...
VAR=$1; VALS=$2; shift 2
for v in $VALS; do echo "== $VAR=$v"; env $VAR=$v timeout 300 python tools/layer_bench.py "$@" 2>&1 | grep b256; done
