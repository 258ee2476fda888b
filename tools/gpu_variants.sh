#!/bin/bash
# usage: gpu_variants.sh "layers" variant...   ("base" = the default library)
L=$1; shift
for v in "$@"; do
  echo "== $v"
  if [ "$v" = base ]; then unset PL_INT8_LIB; else export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_$v.so; fi
  timeout 300 python tools/layer_bench.py $L 2>&1 | grep b256
done
