// fp32_tail.cu -- the un-quantised ops a quantised segmentation model (humanseg_shfv75_0119_quant) keeps between
// and after its int8 convs: channel gather (concat / shuffle_channel / split chains +
// the calib that follows them), conv2d_transpose, elementwise_add, softmax over channels, scale.  All NHWC fp32,
// bandwidth-bound or tiny; no tensor cores.
//   conv2d_transpose   lite/kernels/arm/conv_transpose_compute.cc:1-230 (gemm + col2im); golden deconv_basic
//                      lite/tests/utils/naive_math_impl.h:540-642
//   softmax            lite/kernels/cuda/softmax_compute.cu
//   scale              lite/kernels/cuda/scale_compute.cc
//   elementwise_add    lite/backends/cuda/math/elementwise.cu
//   concat/split/shuffle_channel  lite/backends/arm/math/{concat.h:29-60, split.cc:54-83, shuffle_channel.cc:24-76}
//   calib              lite/kernels/cuda/calib_compute.cu:27-95 (ARM convention: x * (1 / scale), half away, [-127, 127])
#include "common.cuh"

namespace pl {

static constexpr int kGatherMaxSrc = 16;
struct GatherSrc {
  const float* ptr[kGatherMaxSrc];
  int ld[kGatherMaxSrc];
};

// dst[p, j] = srcs[table[j].x][p * ld + table[j].y], optionally quantised on the way out.  One thread per
// (pixel, 4 destination channels): int8 outputs leave as one 32-bit word, fp32 outputs as one float4 when aligned.
__global__ void __launch_bounds__(256)
channel_gather_kernel(const GatherSrc s, const int2* __restrict__ table, void* __restrict__ dst, int ldd, int C,
                      long long pixels, int out_i8, float inv, int round_mode) {
  pdl_launch_dependents();
  pdl_wait();
  const int cq = (C + 3) >> 2;
  const long long total = pixels * cq;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / cq;
    const int c0 = (int)(i - p * cq) * 4;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (c0 + j < C) {
        const int2 t = __ldg(table + c0 + j);
        v[j] = __ldg(s.ptr[t.x] + p * s.ld[t.x] + t.y);
      }
    }
    if (out_i8) {
      int q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q[j] = c0 + j < C ? round_sat_i8(__fmul_rn(inv, v[j]), round_mode) : 0;
      int8_t* o = reinterpret_cast<int8_t*>(dst) + p * ldd + c0;
      if (c0 + 4 <= ldd && (ldd & 3) == 0)
        *reinterpret_cast<uint32_t*>(o) = pack4_i8(q[0], q[1], q[2], q[3]);  // pad channels inside the pitch: 0
      else
        for (int j = 0; j < 4 && c0 + j < C; ++j) o[j] = (int8_t)q[j];
    } else {
      float* o = reinterpret_cast<float*>(dst) + p * ldd + c0;
      if (c0 + 4 <= C && (ldd & 3) == 0)
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
      else
        for (int j = 0; j < 4 && c0 + j < C; ++j) o[j] = v[j];
    }
  }
}

__global__ void __launch_bounds__(256)
add_f32_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ y, long long n, int relu) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = __fadd_rn(__ldg(a + i), __ldg(b + i));
    y[i] = relu ? (v > 0.f ? v : 0.f) : v;
  }
}

// conv2d_transpose, gather form: out[n, p, q, co] = bias[co] + sum over (kh, kw, ci) of x[n, ih, iw, ci] *
// w[ci, co, kh, kw] with ih = (p + pad - kh) / stride when divisible.  Accumulation order is part of the contract
// (kh, kw ascending, then ci ascending, one fmaf each; bias added last) so that the CPU oracle can reproduce it
// bit for bit -- a quantised consumer follows.  One thread per output element, consecutive threads = consecutive co.
__global__ void __launch_bounds__(256)
deconv_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                  void* __restrict__ y, int N, int H, int W, int Ci, int Co, int kh, int kw, int sh, int sw, int pt,
                  int pl_, int P, int Q, int ldy, int act, float act_arg, int out_i8, float inv, int round_mode) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = (long long)N * P * Q * Co;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int co = (int)(i % Co);
    long long t = i / Co;
    const int q = (int)(t % Q);
    t /= Q;
    const int p = (int)(t % P);
    const int n = (int)(t / P);
    float acc = 0.f;
    for (int r = 0; r < kh; ++r) {
      const int th = p + pt - r;
      if (th < 0 || th % sh) continue;
      const int ih = th / sh;
      if (ih >= H) continue;
      for (int s = 0; s < kw; ++s) {
        const int tw = q + pl_ - s;
        if (tw < 0 || tw % sw) continue;
        const int iw = tw / sw;
        if (iw >= W) continue;
        const float* xp = x + (((long long)n * H + ih) * W + iw) * Ci;
        const float* wp = w + ((long long)co * kh + r) * kw + s;  // + ci * Co * kh * kw
        for (int ci = 0; ci < Ci; ++ci) acc = __fmaf_rn(__ldg(xp + ci), __ldg(wp + (long long)ci * Co * kh * kw), acc);
      }
    }
    float v = __fadd_rn(acc, bias ? __ldg(bias + co) : 0.f);
    v = apply_act(v, act, act_arg);
    const long long o = (((long long)n * P + p) * Q + q) * ldy + co;
    if (out_i8)
      reinterpret_cast<int8_t*>(y)[o] = (int8_t)round_sat_i8(__fmul_rn(inv, v), round_mode);
    else
      reinterpret_cast<float*>(y)[o] = v;
  }
}

// softmax over the channels of every pixel (axis = 1 of the model's NCHW view); one thread per pixel
__global__ void __launch_bounds__(256)
softmax_f32_kernel(const float* __restrict__ x, float* __restrict__ y, long long pixels, int C, int ldx, int ldy) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < pixels;
       p += (long long)gridDim.x * blockDim.x) {
    const float* xp = x + p * ldx;
    float m = xp[0];
    for (int c = 1; c < C; ++c) m = fmaxf(m, xp[c]);
    float sum = 0.f;
    for (int c = 0; c < C; ++c) sum += expf(xp[c] - m);
    const float r = 1.f / sum;
    for (int c = 0; c < C; ++c) y[p * ldy + c] = expf(xp[c] - m) * r;
  }
}

__global__ void __launch_bounds__(256)
scale_f32_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, float scale, float bias, int after) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = __ldg(x + i);
    y[i] = after ? __fmaf_rn(v, scale, bias) : __fmul_rn(__fadd_rn(v, bias), scale);
  }
}

static unsigned grid_for(long long n) {
  long long b = (n + 255) / 256;
  if (b > 148 * 16) b = 148 * 16;
  return (unsigned)(b < 1 ? 1 : b);
}

}  // namespace pl

using namespace pl;

extern "C" int pl_channel_gather_f32_nhwc(int n_src, const float* const* srcs, const int* src_ld, const int32_t* table_dev,
                                          void* dst, int ldd, int C, int64_t pixels, int out_dtype, float scale,
                                          int round_mode, pl_stream_t stream) {
  if (n_src <= 0 || n_src > kGatherMaxSrc || !srcs || !src_ld || !table_dev || !dst || C <= 0 || ldd < C || pixels <= 0)
    return PL_ERR_ARG;
  if (out_dtype != PL_OUT_INT8 && out_dtype != PL_OUT_FP32) return PL_ERR_ARG;
  if (out_dtype == PL_OUT_INT8 && !(scale > 0.f)) return PL_ERR_ARG;
  GatherSrc s{};
  for (int i = 0; i < n_src; ++i) {
    if (!srcs[i] || src_ld[i] <= 0) return PL_ERR_ARG;
    s.ptr[i] = srcs[i];
    s.ld[i] = src_ld[i];
  }
  if (out_dtype == PL_OUT_FP32 && (reinterpret_cast<uintptr_t>(dst) & 15) && (ldd & 3) == 0) return PL_ERR_ALIGN;
  if (out_dtype == PL_OUT_INT8 && (reinterpret_cast<uintptr_t>(dst) & 3) && (ldd & 3) == 0) return PL_ERR_ALIGN;
  launch_pdl(channel_gather_kernel, dim3(grid_for(pixels * ((C + 3) / 4))), dim3(256), 0, (cudaStream_t)stream, s,
             reinterpret_cast<const int2*>(table_dev), dst, ldd, C, (long long)pixels, out_dtype == PL_OUT_INT8 ? 1 : 0,
             out_dtype == PL_OUT_INT8 ? 1.f / scale : 0.f, round_mode);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_add_f32(const float* a, const float* b, float* y, int64_t n, int act, pl_stream_t stream) {
  if (!a || !b || !y || n <= 0 || (act != PL_ACT_NONE && act != PL_ACT_RELU)) return PL_ERR_ARG;
  launch_pdl(add_f32_kernel, dim3(grid_for(n)), dim3(256), 0, (cudaStream_t)stream, a, b, y, (long long)n,
             act == PL_ACT_RELU ? 1 : 0);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_conv2d_transpose_f32_nhwc(const float* x, const float* w_iohw, const float* bias, void* y, int N, int H,
                                            int W, int Cin, int Cout, int kh, int kw, int stride_h, int stride_w,
                                            int pad_t, int pad_l, int P, int Q, int ldy, int act, float act_arg,
                                            int out_dtype, float out_scale, int round_mode, pl_stream_t stream) {
  if (!x || !w_iohw || !y || N <= 0 || H <= 0 || W <= 0 || Cin <= 0 || Cout <= 0 || kh <= 0 || kw <= 0 || stride_h <= 0 ||
      stride_w <= 0 || pad_t < 0 || pad_l < 0 || P <= 0 || Q <= 0 || ldy < Cout)
    return PL_ERR_ARG;
  if (out_dtype != PL_OUT_INT8 && out_dtype != PL_OUT_FP32) return PL_ERR_ARG;
  if (out_dtype == PL_OUT_INT8 && !(out_scale > 0.f)) return PL_ERR_ARG;
  launch_pdl(deconv_f32_kernel, dim3(grid_for((long long)N * P * Q * Cout)), dim3(256), 0, (cudaStream_t)stream, x, w_iohw,
             bias, y, N, H, W, Cin, Cout, kh, kw, stride_h, stride_w, pad_t, pad_l, P, Q, ldy, act, act_arg,
             out_dtype == PL_OUT_INT8 ? 1 : 0, out_dtype == PL_OUT_INT8 ? 1.f / out_scale : 0.f, round_mode);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_softmax_f32_nhwc(const float* x, float* y, int64_t pixels, int C, int ldx, int ldy, pl_stream_t stream) {
  if (!x || !y || pixels <= 0 || C <= 0 || ldx < C || ldy < C) return PL_ERR_ARG;
  launch_pdl(softmax_f32_kernel, dim3(grid_for(pixels)), dim3(256), 0, (cudaStream_t)stream, x, y, (long long)pixels, C,
             ldx, ldy);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_scale_f32(const float* x, float* y, int64_t n, float scale, float bias, int bias_after_scale,
                            pl_stream_t stream) {
  if (!x || !y || n <= 0) return PL_ERR_ARG;
  launch_pdl(scale_f32_kernel, dim3(grid_for(n)), dim3(256), 0, (cudaStream_t)stream, x, y, (long long)n, scale, bias,
             bias_after_scale ? 1 : 0);
  ++g_launches;
  return launch_status();
}
