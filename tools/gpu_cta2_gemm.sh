#!/bin/bash
# GEMM sweep variants of the CTA-pair launches
run() {
  echo "== $*"
  env "$@" timeout 600 python tools/gemm_sweep.py 4096x4096x4096 8192x8192x8192 16384x16384x16384 8192x4096x8192 4096x8192x4096 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['shape'], d['out'], d['ms_gemm'], d['tops_gemm'], d['frac_mma_peak'])"
}
run PL_CTA2=0
run PL_CTA2=1 PL_NO_MPAIR=1
run PL_CTA2=1 PL_NO_MPAIR=1 PL_GROUP_M=8
run PL_CTA2=1 PL_NO_MPAIR=1 PL_GROUP_M=4
run PL_CTA2=1
