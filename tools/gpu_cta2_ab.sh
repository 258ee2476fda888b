#!/bin/bash
# A/B of the CTA-pair launches (PL_CTA2 = 0 never / 1 auto / 2 wherever legal): dense layers, GEMM sweep, ResNet-50
mkdir -p gpurun_out
L="r2_1x1_256_64 r2_3x3_64 r3_3x3_128 r3_1x1_128_512_res r3_1x1_512_128 r3_1x1_256_128 r4_3x3_256 r4_1x1_256_1024_res r4_1x1_1024_256 r5_3x3_512 r5_1x1_2048_512 r5_1x1_512_2048_res r3_3x3s2_128 r4_3x3s2_256 m_pw_14_512_512"
for v in ${CTA2_VALUES:-0 1}; do
  echo "== PL_CTA2=$v"
  PL_CTA2=$v timeout 600 python tools/layer_bench.py $L --iters 50 2>&1 | grep -v Warn
done | tee gpurun_out/cta2_layers.txt
run() {
  echo "== $*"
  env "$@" timeout 600 python tools/gemm_sweep.py 2048x2048x2048 4096x4096x4096 8192x8192x8192 16384x16384x16384 8192x4096x8192 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['shape'], d['out'], d['ms_gemm'], d['tops_gemm'], d['frac_mma_peak'])"
}
(run PL_CTA2=0; run PL_CTA2=1; run PL_CTA2=1 PL_CTA2_MPAIR=1) | tee gpurun_out/cta2_gemm.txt
for v in 0 1 0 1; do
  PL_CTA2=$v timeout 600 python bench.py --net resnet50 --steps 30 --warmup 5 --no-cpu-baseline 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PL_CTA2=$v', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
done | tee gpurun_out/cta2_net.txt
