// ref_shim.cc -- C-ABI window onto the REFERENCE's own naive int8 math,
// compiled from the sources where they lie (never copied):
//   /root/reference/lite/tests/utils/naive_math_impl.h
//     conv_basic<int8_t,int>   :350-453   (NCHW activations, OIHW filters)
//     basic_gemm<int8_t,int>   :245-295
//     basic_gemv<int8_t,int>   :297-343
// TEST INFRASTRUCTURE ONLY.  Output goes to oracle/_ref/libref_naive.so (git-
// ignored, shipped to the GPU box by gpurun).  Used to validate oracle.c and as
// the "reference" CPU baseline in bench.py.  -DPADDLE_WITH_MKLML switches on the
// header's own `#pragma omp parallel for` lines (:151,262,311,394).
#include <cstdint>
#include <cstdio>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "lite/tests/utils/naive_math_impl.h"

extern "C" {

int ref_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// int8 x int8 -> int32, no bias, no activation: the exact accumulator.
void ref_conv_basic_i8(const int8_t* din_nchw, int32_t* dout_nchw, int num,
                       int chout, int hout, int wout, int chin, int hin,
                       int win, const int8_t* weights_oihw, int group,
                       int kernel_w, int kernel_h, int stride_w, int stride_h,
                       int dila_w, int dila_h, int pad_w, int pad_h) {
  conv_basic<int8_t, int>(din_nchw, dout_nchw, num, chout, hout, wout, chin,
                          hin, win, weights_oihw, nullptr, group, kernel_w,
                          kernel_h, stride_w, stride_h, dila_w, dila_h, pad_w,
                          pad_h, false, 0);
}

// fp32 golden used by lite/tests/math/conv_int8_compute_test.cc:298-321
void ref_conv_basic_f32(const float* din, float* dout, int num, int chout,
                        int hout, int wout, int chin, int hin, int win,
                        const float* weights, const float* bias, int group,
                        int kernel_w, int kernel_h, int stride_w, int stride_h,
                        int dila_w, int dila_h, int pad_w, int pad_h,
                        int flag_bias, int act_type, float six, float alpha) {
  conv_basic<float, float>(din, dout, num, chout, hout, wout, chin, hin, win,
                           weights, bias, group, kernel_w, kernel_h, stride_w,
                           stride_h, dila_w, dila_h, pad_w, pad_h,
                           flag_bias != 0, act_type, six, alpha);
}

void ref_basic_gemm_i8(int trans_a, int trans_b, int m, int n, int k,
                       const int8_t* a, int lda, const int8_t* b, int ldb,
                       int32_t* c, int ldc) {
  std::memset(c, 0, sizeof(int32_t) * static_cast<size_t>(m) * ldc);
  basic_gemm<int8_t, int>(trans_a != 0, trans_b != 0, m, n, k, 1, a, lda, b,
                          ldb, 0, c, ldc, nullptr, false, false);
}

void ref_basic_gemv_i8(int m, int k, const int8_t* a, const int8_t* b,
                       int32_t* c, int trans_a) {
  std::memset(c, 0, sizeof(int32_t) * static_cast<size_t>(m));
  basic_gemv<int8_t, int>(m, k, a, b, nullptr, c, 1, 0, trans_a != 0, false,
                          0);
}

}  // extern "C"
