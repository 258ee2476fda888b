#!/bin/bash
# check_dropin.sh -- proves the drop-in claim: the kCUDA/kInt8 kernel class sources under
# paddle-lite_b200/lite/kernels/cuda/ compile, byte for byte, against the REFERENCE's own headers
# (lite/core/{kernel,tensor,op_registry,type_system,context}.h, lite/operators/op_params.h, ...), not only
# against this repo's miniature of them.
#
# Only CMake-GENERATED files are supplied: lite/core/version.h is produced from the reference's version.h.in
# exactly as configure_file would.  The one SOURCE edit a maintainer has to make is
# applied to a temporary copy: WITH_INT8_CONFIG in MatMulParam and ActivationParam -- the reference has no int8
# fields there (SURVEY.md 8b "needs int8 fields added").
#
#   tools/check_dropin.sh [/root/reference]      exit 0 = every kernel source passes g++ -fsyntax-only
set -euo pipefail
REF=${1:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
CUDA=${CUDA:-/usr/local/cuda}
[ -d "$REF/lite/core" ] || { echo "no reference tree at $REF"; exit 77; }
TMP=$(mktemp -d)
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$TMP/lite/core" "$TMP/lite/operators"
# 1. generated header (cmake: configure_file(version.h.in version.h))
sed -e 's/@[A-Z_]*@/dropin-check/g' "$REF/lite/core/version.h.in" > "$TMP/lite/core/version.h"
# 2. the documented op_params.h edit: int8 config for matmul and activation
sed -e 's/^struct MatMulParam : ParamBase {$/struct MatMulParam : ParamBase {\n  WITH_INT8_CONFIG/' \
    -e 's/^struct ActivationParam : ParamBase {$/struct ActivationParam : ParamBase {\n  WITH_INT8_CONFIG/' \
    "$REF/lite/operators/op_params.h" > "$TMP/lite/operators/op_params.h"
grep -c "WITH_INT8_CONFIG" "$TMP/lite/operators/op_params.h" > /dev/null
rc=0
for f in conv_compute_int8 gemm_compute_int8 misc_compute_int8; do
  # include order: generated/patched files, then the REFERENCE tree, then this repo (include/pl_int8.h and the
  # kernel classes' own headers lite/kernels/cuda/{conv_compute_int8,int8_common}.h, which the reference lacks)
  if g++ -std=c++14 -fsyntax-only -DLITE_WITH_CUDA -w -I"$TMP" -I"$REF" -I"$ROOT" -I"$ROOT/paddle-lite_b200" \
       -I"$CUDA/include" "$ROOT/paddle-lite_b200/lite/kernels/cuda/$f.cc"; then
    echo "ok   $f.cc compiles against $REF/lite headers"
  else
    echo "FAIL $f.cc"
    rc=1
  fi
done
exit $rc
