#!/bin/bash
# experiments: depthwise with 3 CTAs / SM (variant build), wave quantisation of the res4 / res5 layers (forced BN)
D="dw_112_32_s1 dw_112_64_s2 dw_56_128_s1 dw_14_512_s1"
echo "== dw base"; timeout 300 python tools/layer_bench.py $D --iters 50 2>&1 | grep b256
echo "== dw 3 CTAs/SM"; PL_INT8_LIB=paddle-lite_b200/variants/libpl_dw3.so timeout 300 python tools/layer_bench.py $D --iters 50 2>&1 | grep b256
PL_INT8_LIB=paddle-lite_b200/variants/libpl_dw3.so timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "depthwise or dw" 2>&1 | tail -2
L="r4_3x3_256 r4_1x1_1024_256 r5_3x3_512 r5_1x1_2048_512 r5_1x1_512_2048_res r4_3x3s2_256 r4_1x1_256_1024_res r3_3x3s2_128"
for bn in 0 128 64; do
  echo "== PL_FORCE_BN=$bn"; PL_FORCE_BN=$bn timeout 300 python tools/layer_bench.py $L --iters 50 2>&1 | grep b256
done
echo "== fc shapes"
for bn in 0 32 64 128; do
  echo "-- PL_FORCE_BN=$bn"; PL_FORCE_BN=$bn timeout 300 python tools/gemm_sweep.py 64x1000x1024 256x1000x2048 64x1024x1024 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['shape'], d['out'], d['ms_gemm'], d['tops_gemm'])"
done
echo "== no final wait variant"
for lib in "" paddle-lite_b200/variants/libpl_nofw.so "" paddle-lite_b200/variants/libpl_nofw.so; do
  PL_INT8_LIB=$lib timeout 300 python tools/conv1_latency.py 2>/dev/null | cut -c1-120
  for net in mobilenet_v1 resnet50; do
  PL_INT8_LIB=$lib timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('lib=$lib', '$net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
  done
done
PL_INT8_LIB=paddle-lite_b200/variants/libpl_nofw.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_nets.py -q -x 2>&1 | tail -2
