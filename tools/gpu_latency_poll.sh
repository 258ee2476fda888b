#!/bin/bash
export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_poll.so
for f in 0 256 31 287; do
  echo -n "poll build PL_DBG_FLAGS=$f: "
  PL_DBG_FLAGS=$f timeout 120 python tools/conv1_latency.py 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print(d['us_per_launch'])"
done
for f in 0 256; do
  for net in mobilenet_v1 resnet50; do
  PL_DBG_FLAGS=$f timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PL_DBG_FLAGS=$f $net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
  done
done
export PL_INT8_LIB=$PWD/paddle-lite_b200/variants/libpl_dbg.so
PL_DBG_TRACE_ON=1 PL_DBG_TRACE=gpurun_out/trace_conv1.txt timeout 120 python tools/layer_bench.py r2_3x3_64 --batch 1 --iters 1 --nsets 1 > /dev/null 2>&1
python tools/trace_timeline.py gpurun_out/trace_conv1.txt 0 1
cat gpurun_out/trace_conv1.txt | head -3
