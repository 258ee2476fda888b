/*
 * oracle.c -- CPU restatement of Paddle-Lite's int8 conv / GEMM hot path.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Built with gcc:
 *   gcc -O2 -std=c11 -fPIC -shared -ffp-contract=off -fopenmp oracle.c -lm
 * -ffp-contract=off matters: every fused multiply-add below is an explicit
 * fmaf(); no other expression may be contracted (SURVEY.md 8a numerical facts).
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* bench.py --impl reference under torchrun inherits OMP_NUM_THREADS=1: the CPU arm sets its
 * thread count explicitly */
void orc_set_num_threads(int n) {
  if (n > 0) omp_set_num_threads(n);
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/*
 * concat along the channel axis of NCHW tensors: lite/backends/arm/math/concat.h:29-60
 * (concat_func, axis = 1): per input, per image, one memcpy of C_k * H * W elements
 * to the running channel offset.  inner_bytes = H * W * sizeof(T).
 */
int orc_concat_nchw(int n_in, const void* const* xs, const int* cs, void* y,
                    int N, int64_t inner_bytes) {
  int64_t ctot = 0, off = 0;
  for (int k = 0; k < n_in; ++k) ctot += cs[k];
  for (int k = 0; k < n_in; ++k) {
    const char* src = (const char*)xs[k];
    char* dst = (char*)y + off * inner_bytes;
    const int64_t in_sum = (int64_t)cs[k] * inner_bytes;
    for (int n = 0; n < N; ++n) {
      memcpy(dst, src, (size_t)in_sum);
      dst += ctot * inner_bytes;
      src += in_sum;
    }
    off += cs[k];
  }
  return 0;
}

/* split along the channel axis: lite/backends/arm/math/split.cc:54-83 (axis = 1) */
int orc_split_nchw(const void* x, int n_out, void* const* ys, const int* cs,
                   int N, int64_t inner_bytes) {
  int64_t ctot = 0, off = 0;
  for (int k = 0; k < n_out; ++k) ctot += cs[k];
  for (int k = 0; k < n_out; ++k) {
    const char* src = (const char*)x + off * inner_bytes;
    char* dst = (char*)ys[k];
    const int64_t out_after = (int64_t)cs[k] * inner_bytes;
    for (int n = 0; n < N; ++n) {
      memcpy(dst, src, (size_t)out_after);
      src += ctot * inner_bytes;
      dst += out_after;
    }
    off += cs[k];
  }
  return 0;
}

/*
 * shuffle_channel: lite/backends/arm/math/shuffle_channel.cc:24-76 and the baseline of
 * lite/tests/kernels/shuffle_channel_compute_test.cc:38-66:
 *   out channel (j * group + i) = in channel (i * (C / group) + j)
 */
int orc_shuffle_channel_nchw(const void* x, void* y, int group, int N, int C,
                             int64_t spatial_bytes) {
  if (group <= 0 || C % group) return -1;
  const int group_col = C / group;
  for (int n = 0; n < N; ++n) {
    const char* in = (const char*)x + (int64_t)n * C * spatial_bytes;
    char* out = (char*)y + (int64_t)n * C * spatial_bytes;
    for (int i = 0; i < group; ++i)
      for (int j = 0; j < group_col; ++j)
        memcpy(out + (int64_t)(j * group + i) * spatial_bytes,
               in + (int64_t)(i * group_col + j) * spatial_bytes, (size_t)spatial_bytes);
  }
  return 0;
}

/*
 * relu on int8 with a symmetric scale: max(x, 0) commutes with the quantiser, so it is
 * lite/kernels/cuda/relu_compute.cu:23-33 (ReluKernel: x >= 0 ? x : 0) on the integers;
 * hi < 127 clips like relu6 at the caller's quantised 6 / scale.
 */
int orc_relu_i8(const int8_t* x, int8_t* y, int64_t n, int hi) {
  for (int64_t i = 0; i < n; ++i) {
    int v = x[i] > 0 ? x[i] : 0;
    y[i] = (int8_t)(v > hi ? hi : v);
  }
  return 0;
}

/* lite/operators/conv_op.cc:42-53 ConvOutputSize */
int orc_conv_out_size(int in, int k, int dil, int pad0, int pad1, int stride) {
  const int dkernel = dil * (k - 1) + 1;
  return (in + (pad0 + pad1) - dkernel) / stride + 1;
}

/*
 * Scale folding.
 *  fp32-out: s[c] = w_scale[c] * in_scale           (conv_gemmlike.cc:209-226)
 *  int8-out: s[c] = w_scale[c] * in_scale / out_scale, bias[c] /= out_scale,
 *            relu6 six /= out_scale                  (conv_gemmlike.cc:228-264)
 *  a single weight scale is replicated to n_out      (conv_gemmlike.cc:217-221)
 * cudnn_conv.cc:352-372 computes the same products (it requires n == n_out).
 */
int orc_fold_scales(const float* w_scale, int n_w_scale, int n_out,
                    float in_scale, float out_scale, int out_dtype,
                    const float* bias, float* scale_folded, float* bias_folded,
                    int act, float act_arg, float* act_arg_folded) {
  if (n_w_scale != 1 && n_w_scale != n_out) return -1;
  for (int i = 0; i < n_out; ++i) {
    float ws = w_scale[n_w_scale == 1 ? 0 : i];
    if (out_dtype == ORC_OUT_INT8) {
      scale_folded[i] = ws * in_scale / out_scale;
      bias_folded[i] = bias ? bias[i] / out_scale : 0.f;
    } else {
      scale_folded[i] = ws * in_scale;
      bias_folded[i] = bias ? bias[i] : 0.f;
    }
  }
  if (act_arg_folded) {
    *act_arg_folded = act_arg;
    if (act == ORC_ACT_RELU6 && out_dtype == ORC_OUT_INT8)
      *act_arg_folded = act_arg / out_scale;
  }
  return 0;
}

/*
 * Round + saturate one float to int8.
 *  ARM  : saturate_cast<int8_t>(roundf(v)) then "< -127 -> -127"
 *         (gemv_arm_int8.cc:122-124, saturate.h:143-147): round half away from
 *         zero, clamp to [-127, 127].
 *  CUDA : clamp to [-128,127] then __float2int_rn (half to even)
 *         (lite/backends/cuda/math/utils.h:61-66).
 */
int8_t orc_round_sat_i8(float v, int round_mode) {
  if (round_mode == ORC_ROUND_CUDA_LEGACY) {
    v = fmaxf(v, -128.f);
    v = fminf(v, 127.f);
    return (int8_t)(int)nearbyintf(v); /* default FE_TONEAREST = half-even */
  }
  if (v != v) return 0; /* NaN: never produced by finite scales; pin to 0 */
  float r = roundf(v);
  if (r > 127.f) r = 127.f;
  if (r < -127.f) r = -127.f;
  return (int8_t)(int)r;
}

/*
 * Epilogue on one accumulator, returning the pre-rounding float.
 *   v = fma((float)acc, scale, bias)        gemm_prepacked_int8.cc:664-700 (fmla)
 *   relu  : v > 0 ? v : 0                   gemv_arm_int8.cc:111-112
 *   relu6 : v > 0 ? (v > six ? six : v) : 0 gemv_arm_int8.cc:113-114
 *   leaky : v > 0 ? v : v * alpha           gemv_arm_int8.cc:115-116
 * CUDA-legacy (activation.cu:91-181): same fma (nvcc contracts a*b+c), then
 *   vec4 path (C % 4 == 0): fmaxf(v*alpha, v); scalar path: v>0 ? v : v*alpha,
 *   alpha = 0 for relu.  relu6 does not exist on that path.
 */
float orc_epilogue_f32(int32_t acc, float scale, float bias, int act,
                       float act_arg, int round_mode, int vec4_path) {
  float v = fmaf((float)acc, scale, bias);
  if (round_mode == ORC_ROUND_CUDA_LEGACY) {
    if (act == ORC_ACT_RELU || act == ORC_ACT_LEAKY) {
      float alpha = (act == ORC_ACT_RELU) ? 0.f : act_arg;
      if (vec4_path)
        v = fmaxf(v * alpha, v);
      else
        v = v > 0.f ? v : v * alpha;
    } else if (act == ORC_ACT_RELU6) {
      v = v > 0.f ? (v > act_arg ? act_arg : v) : 0.f;
    }
    return v;
  }
  switch (act) {
    case ORC_ACT_RELU:
      v = v > 0.f ? v : 0.f;
      break;
    case ORC_ACT_RELU6:
      v = v > 0.f ? (v > act_arg ? act_arg : v) : 0.f;
      break;
    case ORC_ACT_LEAKY:
      v = v > 0.f ? v : v * act_arg;
      break;
    default:
      break;
  }
  return v;
}

int orc_epilogue(const int32_t* acc, int64_t M, int n_out, const float* scale,
                 const float* bias, int act, float act_arg, int out_dtype,
                 int round_mode, void* y) {
  if (out_dtype == ORC_OUT_INT32) {
    memcpy(y, acc, (size_t)M * n_out * sizeof(int32_t));
    return 0;
  }
  const int vec4 = (n_out % 4 == 0);
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    for (int c = 0; c < n_out; ++c) {
      float v = orc_epilogue_f32(acc[m * n_out + c], scale[c],
                                 bias ? bias[c] : 0.f, act, act_arg,
                                 round_mode, vec4);
      if (out_dtype == ORC_OUT_FP32)
        ((float*)y)[m * n_out + c] = v;
      else
        ((int8_t*)y)[m * n_out + c] = orc_round_sat_i8(v, round_mode);
    }
  }
  return 0;
}

/*
 * Epilogue with a fused int8 residual (ConvParam::residualData,
 * lite/operators/op_params.h:446-502; the reference fuses conv + elementwise_add
 * only for fp32, lite/core/mir/fusion/conv_elementwise_fuser.cc).  PARITY
 * UNPINNED: defined here by analogy with orc_add_i8:
 *   v = fmaf((float)acc, scale, bias); v = fmaf((float)res, res_scale, v); act; round
 */
int orc_epilogue_res(const int32_t* acc, int64_t M, int n_out, const float* scale,
                     const float* bias, const int8_t* res, int ld_res,
                     float res_scale, int act, float act_arg, int out_dtype,
                     int round_mode, void* y) {
  const int vec4 = (n_out % 4 == 0);
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < M; ++m) {
    for (int c = 0; c < n_out; ++c) {
      float v = fmaf((float)acc[m * n_out + c], scale[c], bias ? bias[c] : 0.f);
      v = fmaf((float)res[m * ld_res + c], res_scale, v);
      if (round_mode == ORC_ROUND_CUDA_LEGACY) {
        if (act == ORC_ACT_RELU || act == ORC_ACT_LEAKY) {
          float alpha = (act == ORC_ACT_RELU) ? 0.f : act_arg;
          v = vec4 ? fmaxf(v * alpha, v) : (v > 0.f ? v : v * alpha);
        } else if (act == ORC_ACT_RELU6) {
          v = v > 0.f ? (v > act_arg ? act_arg : v) : 0.f;
        }
      } else if (act == ORC_ACT_RELU) {
        v = v > 0.f ? v : 0.f;
      } else if (act == ORC_ACT_RELU6) {
        v = v > 0.f ? (v > act_arg ? act_arg : v) : 0.f;
      } else if (act == ORC_ACT_LEAKY) {
        v = v > 0.f ? v : v * act_arg;
      }
      if (out_dtype == ORC_OUT_FP32)
        ((float*)y)[m * n_out + c] = v;
      else
        ((int8_t*)y)[m * n_out + c] = orc_round_sat_i8(v, round_mode);
    }
  }
  return 0;
}

/*
 * Exact int32 convolution, NHWC activations, OHWI filters, groups, dilation,
 * top/left padding (bottom/right only enter through P,Q).  Same loop nest and
 * index arithmetic as conv_basic<int8_t,int> (naive_math_impl.h:350-453) with
 * the layouts changed from NCHW/OIHW; integer addition is associative so the
 * accumulator is identical.
 */
int orc_conv2d_i8_acc(const int8_t* x, const int8_t* w, int32_t* acc, int N,
                      int H, int W, int C, int K, int R, int S, int sh, int sw,
                      int pt, int pl, int dh, int dw, int groups, int P,
                      int Q) {
  if (groups <= 0 || C % groups || K % groups) return -1;
  const int cg = C / groups, kg = K / groups;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int p = 0; p < P; ++p) {
      for (int q = 0; q < Q; ++q) {
        int32_t* out = acc + (((int64_t)n * P + p) * Q + q) * K;
        for (int k = 0; k < K; ++k) {
          const int g = k / kg;
          int32_t sum = 0;
          for (int r = 0; r < R; ++r) {
            const int ih = p * sh - pt + r * dh;
            if (ih < 0 || ih >= H) continue;
            for (int s = 0; s < S; ++s) {
              const int iw = q * sw - pl + s * dw;
              if (iw < 0 || iw >= W) continue;
              const int8_t* xp =
                  x + (((int64_t)n * H + ih) * W + iw) * C + (int64_t)g * cg;
              const int8_t* wp = w + (((int64_t)k * R + r) * S + s) * cg;
              for (int c = 0; c < cg; ++c) sum += (int32_t)xp[c] * (int32_t)wp[c];
            }
          }
          out[k] = sum;
        }
      }
    }
  }
  return 0;
}

/* basic_gemm<int8_t,int> (naive_math_impl.h:245-295) with alpha=1, beta=0. */
int orc_gemm_i8_acc(const int8_t* A, const int8_t* B, int32_t* acc, int M,
                    int N, int K, int trans_a, int b_is_nk) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < M; ++i) {
    for (int j = 0; j < N; ++j) {
      int32_t sum = 0;
      for (int l = 0; l < K; ++l) {
        int8_t av = trans_a ? A[(int64_t)l * M + i] : A[(int64_t)i * K + l];
        int8_t bv = b_is_nk ? B[(int64_t)j * K + l] : B[(int64_t)l * N + j];
        sum += (int32_t)av * (int32_t)bv;
      }
      acc[(int64_t)i * N + j] = sum;
    }
  }
  return 0;
}

int orc_conv2d_i8_nhwc(const int8_t* x, const int8_t* w, const float* bias_f,
                       const float* scale_f, void* y, int N, int H, int W,
                       int C, int K, int R, int S, int sh, int sw, int pt,
                       int pb, int pl, int pr, int dh, int dw, int groups,
                       int act, float act_arg, int out_dtype, int round_mode) {
  const int P = orc_conv_out_size(H, R, dh, pt, pb, sh);
  const int Q = orc_conv_out_size(W, S, dw, pl, pr, sw);
  if (P <= 0 || Q <= 0) return -1;
  const int64_t M = (int64_t)N * P * Q;
  int32_t* acc = (int32_t*)malloc((size_t)M * K * sizeof(int32_t));
  if (!acc) return -2;
  int rc = orc_conv2d_i8_acc(x, w, acc, N, H, W, C, K, R, S, sh, sw, pt, pl,
                             dh, dw, groups, P, Q);
  if (!rc)
    rc = orc_epilogue(acc, M, K, scale_f, bias_f, act, act_arg, out_dtype,
                      round_mode, y);
  free(acc);
  return rc;
}

int orc_gemm_i8(const int8_t* A, const int8_t* B, const float* bias_f,
                const float* scale_f, void* y, int M, int N, int K,
                int trans_a, int b_is_nk, int act, float act_arg,
                int out_dtype, int round_mode) {
  int32_t* acc = (int32_t*)malloc((size_t)M * N * sizeof(int32_t));
  if (!acc) return -2;
  int rc = orc_gemm_i8_acc(A, B, acc, M, N, K, trans_a, b_is_nk);
  if (!rc)
    rc = orc_epilogue(acc, M, N, scale_f, bias_f, act, act_arg, out_dtype,
                      round_mode, y);
  free(acc);
  return rc;
}

/*
 * calib fp32 -> int8.
 *  ARM  (type_trans.cc:34-187): inv = 1.f / scale; q = sat(roundf(inv * x)),
 *        floor at -127.  Golden: calib_compute_test.cc:54-69.
 *  CUDA (calib_compute.cu:27-36): from_float<int8_t>(x / scale): clamp
 *        [-128,127], round half-even.  Golden: calib_compute_cuda_test.cc:35-46
 *        (roundf there; differs from the kernel only on exact .5 ties).
 */
int orc_quantize_f32_i8(const float* x, int8_t* y, int64_t n, float scale,
                        int round_mode) {
  const float inv = 1.f / scale;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    float v = (round_mode == ORC_ROUND_CUDA_LEGACY) ? x[i] / scale : inv * x[i];
    y[i] = orc_round_sat_i8(v, round_mode);
  }
  return 0;
}

/* calib int8 -> fp32: q * scale (type_trans.cc:268-..., calib_compute.cu:38-46) */
int orc_dequantize_i8_f32(const int8_t* x, float* y, int64_t n, float scale) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) y[i] = (float)x[i] * scale;
  return 0;
}

/* layout: [N,C,HW] <-> [N,HW,C] batched transpose (layout_compute.cc:41-94) */
int orc_nchw_to_nhwc(const void* x, void* y, int N, int C, int H, int W,
                     int eb) {
  const int64_t HW = (int64_t)H * W;
  const char* xi = (const char*)x;
  char* yo = (char*)y;
#pragma omp parallel for schedule(static)
  for (int64_t nhw = 0; nhw < (int64_t)N * HW; ++nhw) {
    const int64_t n = nhw / HW, hw = nhw % HW;
    for (int c = 0; c < C; ++c)
      memcpy(yo + (nhw * C + c) * eb, xi + ((n * C + c) * HW + hw) * eb, eb);
  }
  return 0;
}

int orc_nhwc_to_nchw(const void* x, void* y, int N, int C, int H, int W,
                     int eb) {
  const int64_t HW = (int64_t)H * W;
  const char* xi = (const char*)x;
  char* yo = (char*)y;
#pragma omp parallel for schedule(static)
  for (int64_t nhw = 0; nhw < (int64_t)N * HW; ++nhw) {
    const int64_t n = nhw / HW, hw = nhw % HW;
    for (int c = 0; c < C; ++c)
      memcpy(yo + ((n * C + c) * HW + hw) * eb, xi + (nhw * C + c) * eb, eb);
  }
  return 0;
}

/*
 * int8 NHWC pooling.  PARITY UNPINNED: the reference has no int8 pool kernel on
 * any target; window/exclusive semantics follow the fp32 kernels
 * (lite/kernels/cuda/pool_compute.cu:28-130): window clipped to the image,
 * exclusive -> divide by the clipped count, else by kh*kw.
 *   max : exact in int8 (then optional requant when rescale != 1)
 *   avg : v = ((float)sum / (float)count) * rescale
 * rescale = in_scale / out_scale (int8 out) or in_scale (fp32 out).
 */
int orc_pool2d_i8_nhwc(const int8_t* x, void* y, int N, int H, int W, int C,
                       int kh, int kw, int sh, int sw, int pt, int pl, int P,
                       int Q, int pool_type, int exclusive, float rescale,
                       int out_dtype, int round_mode) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int p = 0; p < P; ++p) {
      for (int q = 0; q < Q; ++q) {
        int h0 = p * sh - pt, w0 = q * sw - pl;
        int h1 = h0 + kh < H ? h0 + kh : H, w1 = w0 + kw < W ? w0 + kw : W;
        if (h0 < 0) h0 = 0;
        if (w0 < 0) w0 = 0;
        const int cnt = exclusive ? (h1 - h0) * (w1 - w0) : kh * kw;
        for (int c = 0; c < C; ++c) {
          int32_t sum = 0, mx = -128;
          for (int h = h0; h < h1; ++h)
            for (int w = w0; w < w1; ++w) {
              int32_t v = x[(((int64_t)n * H + h) * W + w) * C + c];
              sum += v;
              if (v > mx) mx = v;
            }
          float v;
          if (pool_type == ORC_POOL_MAX)
            v = (float)mx * rescale;
          else
            v = ((float)sum / (float)cnt) * rescale;
          const int64_t o = (((int64_t)n * P + p) * Q + q) * C + c;
          if (out_dtype == ORC_OUT_FP32)
            ((float*)y)[o] = v;
          else
            ((int8_t*)y)[o] = orc_round_sat_i8(v, round_mode);
        }
      }
    }
  }
  return 0;
}

/*
 * adaptive int8 NHWC pooling (PoolParam::adaptive, ksize = output size): window
 * [floor(i * in / out), ceil((i + 1) * in / out)) per axis, avg divides by the window
 * size -- lite/kernels/cuda/pool_compute.cu:69-106 (max), :150-187 (avg).  Value rules
 * as orc_pool2d_i8_nhwc.
 */
int orc_adaptive_pool2d_i8_nhwc(const int8_t* x, void* y, int N, int H, int W, int C, int P, int Q, int pool_type,
                                float rescale, int out_dtype, int round_mode) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int p = 0; p < P; ++p) {
      for (int q = 0; q < Q; ++q) {
        const int h0 = (p * H) / P, h1 = ((p + 1) * H + P - 1) / P;
        const int w0 = (q * W) / Q, w1 = ((q + 1) * W + Q - 1) / Q;
        const int cnt = (h1 - h0) * (w1 - w0);
        for (int c = 0; c < C; ++c) {
          int32_t sum = 0, mx = -128;
          for (int h = h0; h < h1; ++h)
            for (int w = w0; w < w1; ++w) {
              int32_t v = x[(((int64_t)n * H + h) * W + w) * C + c];
              sum += v;
              if (v > mx) mx = v;
            }
          const float v = pool_type == ORC_POOL_MAX ? (float)mx * rescale : ((float)sum / (float)cnt) * rescale;
          const int64_t o = (((int64_t)n * P + p) * Q + q) * C + c;
          if (out_dtype == ORC_OUT_FP32)
            ((float*)y)[o] = v;
          else
            ((int8_t*)y)[o] = orc_round_sat_i8(v, round_mode);
        }
      }
    }
  }
  return 0;
}

/*
 * conv2d_transpose fp32 NHWC, gather form with a FIXED accumulation order (kh, kw ascending, then ci ascending,
 * one fmaf each, bias added last): the value is what lite/tests/utils/naive_math_impl.h:540-642 (deconv_basic:
 * gemm + col2im) computes up to fp32 summation order -- tests/test_oracle.py checks the two within 1e-5 -- and
 * the order is pinned because a quantised consumer follows this op in the shipped segmentation model.
 * w: [Cin, Cout, kh, kw] (the reference's layout), dilation 1, groups 1.  out_scale > 0: int8 output through the
 * ARM calib (x * (1 / scale), half away, [-127, 127]).
 */
int orc_conv2d_transpose_f32_nhwc(const float* x, const float* w, const float* bias, void* y, int N, int H, int W,
                                  int Ci, int Co, int kh, int kw, int sh, int sw, int pt, int pl, int P, int Q,
                                  int relu, float out_scale) {
  const float inv = out_scale > 0.f ? 1.f / out_scale : 0.f;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n)
    for (int p = 0; p < P; ++p)
      for (int q = 0; q < Q; ++q)
        for (int co = 0; co < Co; ++co) {
          float acc = 0.f;
          for (int r = 0; r < kh; ++r) {
            const int th = p + pt - r;
            if (th < 0 || th % sh || th / sh >= H) continue;
            for (int s = 0; s < kw; ++s) {
              const int tw = q + pl - s;
              if (tw < 0 || tw % sw || tw / sw >= W) continue;
              const float* xp = x + (((int64_t)n * H + th / sh) * W + tw / sw) * Ci;
              for (int ci = 0; ci < Ci; ++ci)
                acc = fmaf(xp[ci], w[(((int64_t)ci * Co + co) * kh + r) * kw + s], acc);
            }
          }
          float v = acc + (bias ? bias[co] : 0.f);
          if (relu) v = v > 0.f ? v : 0.f;
          const int64_t o = (((int64_t)n * P + p) * Q + q) * Co + co;
          if (out_scale > 0.f)
            ((int8_t*)y)[o] = orc_round_sat_i8(inv * v, ORC_ROUND_ARM);
          else
            ((float*)y)[o] = v;
        }
  return 0;
}

/*
 * int8 residual add + act + requant.  PARITY UNPINNED (the
 * reference's elementwise_add_int8_kernel takes fp32 inputs and is dead code).
 *   v = fma((float)a, sa, (float)b * sb), sa = a_scale/out_scale, sb likewise
 */
int orc_add_i8(const int8_t* a, const int8_t* b, void* y, int64_t n, float sa,
               float sb, int act, float act_arg, int out_dtype,
               int round_mode) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    float v = fmaf((float)a[i], sa, (float)b[i] * sb);
    if (act == ORC_ACT_RELU)
      v = v > 0.f ? v : 0.f;
    else if (act == ORC_ACT_RELU6)
      v = v > 0.f ? (v > act_arg ? act_arg : v) : 0.f;
    else if (act == ORC_ACT_LEAKY)
      v = v > 0.f ? v : v * act_arg;
    if (out_dtype == ORC_OUT_FP32)
      ((float*)y)[i] = v;
    else
      ((int8_t*)y)[i] = orc_round_sat_i8(v, round_mode);
  }
  return 0;
}
