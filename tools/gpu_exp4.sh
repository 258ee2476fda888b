#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_cta2.py tests/test_gpu_parity.py tests/test_gpu_nets.py -q -x 2>&1 | tail -2
L="stem7x7_s2d stem3x3_s2d r2_3x3_64 r2_1x1_64_64 r2_1x1_256_64 m_pw_112_32_64 r2_1x1_64_256_res r3_3x3_128 r3_1x1_128_512_res"
run() { echo "== $*"; env "$@" timeout 300 python tools/layer_bench.py $L --iters 50 2>&1 | grep b256 | cut -c1-110; }
run PL_STORE_DEFER=0
run PL_STORE_DEFER=1
run PL_STORE_DEFER=2
run PL_STORE_DEFER=3 PL_IGEMM_OUT_BUFS=6
run PL_STORE_DEFER=4 PL_IGEMM_OUT_BUFS=8
run PL_STORE_DEFER=2 PL_IGEMM_OUT_BUFS=8
for v in "PL_STORE_DEFER=0" "PL_STORE_DEFER=-1" "PL_STORE_DEFER=1" "PL_STORE_DEFER=4 PL_IGEMM_OUT_BUFS=8"; do
 for net in mobilenet_v1 resnet50 humanseg; do
  env $v timeout 600 python bench.py --net $net --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/ab_err.log | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v $net', 'ms/step %.4f'%l['ms_per_step'], 'img/s %.0f'%l['value'])"
 done
done
