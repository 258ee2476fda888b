// ref_shim.cc -- C-ABI window onto the REFERENCE's own naive int8 math,
// compiled from the sources where they lie (never copied):
//   /root/reference/lite/tests/utils/naive_math_impl.h
//     conv_basic<int8_t,int>   :350-453   (NCHW activations, OIHW filters)
//     basic_gemm<int8_t,int>   :245-295
//     basic_gemv<int8_t,int>   :297-343
//     deconv_basic<float,float> :540-642  (conv2d_transpose golden, fp32)
//   /root/reference/lite/backends/arm/math/saturate.h
//     saturate_cast<int8_t>(float) / (int)  :130-147  (round half away + clamp:
//     the last step of every ARM int8 epilogue, conv_block_utils.h:3201-3224)
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
#include "lite/backends/arm/math/saturate.h"

extern "C" {

int ref_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void ref_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// the reference's own rounding + saturation object code (saturate.h:144-147: roundf, then the int
// overload's clamp to [-128, 127]); the -127 floor is the caller's (conv_block_utils.h:3210 etc.)
int8_t ref_saturate_cast_i8_f32(float v) {
  return paddle::lite::arm::math::saturate_cast<int8_t>(v);
}
int8_t ref_saturate_cast_i8_i32(int v) {
  return paddle::lite::arm::math::saturate_cast<int8_t>(v);
}
void ref_saturate_cast_i8_f32_n(const float* v, int8_t* out, long long n) {
  for (long long i = 0; i < n; ++i) out[i] = paddle::lite::arm::math::saturate_cast<int8_t>(v[i]);
}

// conv2d_transpose golden (fp32, NCHW activations, weights [chin, chout/group, kh, kw])
void ref_deconv_basic_f32(const float* din, float* dout, int num, int chout, int hout, int wout, int chin,
                          int hin, int win, const float* weights, const float* bias, int group,
                          int kernel_w, int kernel_h, int stride_w, int stride_h, int dila_w, int dila_h,
                          int pad_w0, int pad_w1, int pad_h0, int pad_h1, int flag_bias, int flag_relu) {
  deconv_basic<float, float>(din, dout, num, chout, hout, wout, chin, hin, win, weights, bias, group,
                             kernel_w, kernel_h, stride_w, stride_h, dila_w, dila_h, pad_w0, pad_w1,
                             pad_h0, pad_h1, flag_bias != 0, flag_relu != 0);
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
