#!/bin/bash
# Experimental build of the C-ABI library with extra -D flags:
#   tools/build_variant.sh NAME "-DPL_EXP_X ..."  ->  paddle-lite_b200/variants/libpl_NAME.so
# run with PL_INT8_LIB=paddle-lite_b200/variants/libpl_NAME.so
set -e
cd "$(dirname "$0")/../paddle-lite_b200/csrc"
NAME=$1; FLAGS=$2
mkdir -p ../variants /tmp/plv_$NAME
for f in igemm_i8 gemv_i8 dwconv_i8 elementwise_i8 channel_ops_i8 fp32_tail capi; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC $FLAGS -c $f.cu -o /tmp/plv_$NAME/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libpl_$NAME.so /tmp/plv_$NAME/*.o -lcudart
ls -la ../variants/libpl_$NAME.so
