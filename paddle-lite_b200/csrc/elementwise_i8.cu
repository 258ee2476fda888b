// elementwise_i8.cu -- the HBM-bound companions of the int8 hot path: calib
// (quantise / dequantise), NCHW<->NHWC layout, int8 pooling, residual add.
// All of them move each byte once with 16-byte vectors; no tensor cores.
//
// Reference semantics (chenjiaoAngel/Paddle-Lite):
//   quantise    ARM  q = sat(roundf(x * (1/scale))), floor -127
//                    lite/backends/arm/math/type_trans.cc:34-187
//               CUDA q = from_float<int8_t>(x / scale)
//                    lite/kernels/cuda/calib_compute.cu:27-36, utils.h:61-66
//   dequantise  x = q * scale   lite/kernels/cuda/calib_compute.cu:38-46
//   layout      [N,C,HW] <-> [N,HW,C]  lite/backends/cuda/math/transpose.cu:31-184
//   pool / add  no int8 kernel in the reference;
//               window rules of lite/kernels/cuda/pool_compute.cu:28-130
#include <cstdlib>
#include <cuda_fp16.h>

#include "common.cuh"

namespace pl {

static inline int grid_for(long long items, int threads, int per_sm = 8) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long g = (items + threads - 1) / threads;
  const long long cap = (long long)sms * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

__device__ __forceinline__ float quant_pre(float x, float scale, float inv, int round_mode) {
  return round_mode == PL_ROUND_CUDA_LEGACY ? __fdiv_rn(x, scale) : __fmul_rn(inv, x);
}

// ---------------------------------------------------------------- quantise
__global__ void __launch_bounds__(256)
quantize_kernel(const float* __restrict__ x, int8_t* __restrict__ y, long long n, float scale,
                float inv, int round_mode, int vec) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const long long n16 = n >> 4;
    for (; i < n16; i += stride) {
      const float4* xp = reinterpret_cast<const float4*>(x) + i * 4;
      float4 f[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) f[k] = __ldg(xp + k);
      uint32_t o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        o[k] = pack4_i8(round_sat_i8(quant_pre(f[k].x, scale, inv, round_mode), round_mode),
                        round_sat_i8(quant_pre(f[k].y, scale, inv, round_mode), round_mode),
                        round_sat_i8(quant_pre(f[k].z, scale, inv, round_mode), round_mode),
                        round_sat_i8(quant_pre(f[k].w, scale, inv, round_mode), round_mode));
      reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    i = (n16 << 4) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  }
  for (; i < n; i += stride)
    y[i] = (int8_t)round_sat_i8(quant_pre(x[i], scale, inv, round_mode), round_mode);
}

// -------------------------------------------------------------- dequantise
__global__ void __launch_bounds__(256)
dequantize_kernel(const int8_t* __restrict__ x, float* __restrict__ y, long long n, float scale,
                  int vec) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const long long n16 = n >> 4;
    for (; i < n16; i += stride) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(x) + i);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      float4* yp = reinterpret_cast<float4*>(y) + i * 4;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float4 f;
        f.x = __fmul_rn((float)(int8_t)(w[k] & 0xff), scale);
        f.y = __fmul_rn((float)(int8_t)((w[k] >> 8) & 0xff), scale);
        f.z = __fmul_rn((float)(int8_t)((w[k] >> 16) & 0xff), scale);
        f.w = __fmul_rn((float)(int8_t)(w[k] >> 24), scale);
        yp[k] = f;
      }
    }
    i = (n16 << 4) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  }
  for (; i < n; i += stride) y[i] = __fmul_rn((float)x[i], scale);
}

// -------------------------------------------- layout: batched 2-D transposes
// src [B][rows][cols] (pitch ld_src) -> dst [B][cols][rows] (pitch ld_dst);
// 64x64 byte tiles through shared memory.  T is the element type.
template <typename T>
__global__ void __launch_bounds__(256)
transpose_kernel(const T* __restrict__ src, T* __restrict__ dst, int rows, int cols,
                 long long ld_src, long long ld_dst, long long bs_src, long long bs_dst) {
  __shared__ T tile[64][65];
  const int b = blockIdx.z;
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const T* s = src + (long long)b * bs_src;
  T* d = dst + (long long)b * bs_dst;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
#pragma unroll 4
  for (int i = ty; i < 64; i += 4) {
    const int r = r0 + i, c = c0 + tx;
    if (r < rows && c < cols) tile[i][tx] = s[(long long)r * ld_src + c];
  }
  __syncthreads();
#pragma unroll 4
  for (int i = ty; i < 64; i += 4) {
    const int c = c0 + i, r = r0 + tx;
    if (r < rows && c < cols) d[(long long)c * ld_dst + r] = tile[tx][i];
  }
}

// zero the pad channels [C, ld) of an NHWC buffer
__global__ void zero_pad_kernel(int8_t* y, long long npix, int C, int ld, int eb) {
  const int padb = (ld - C) * eb;
  const long long total = npix * padb;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long px = i / padb;
    const int o = (int)(i - px * padb);
    y[(px * ld + C) * eb + o] = 0;
  }
}

// fused entry of the int8 region: fp32 NCHW -> int8 NHWC(ld), pad channels = 0.
// One block stages a [C x 64 pixels] slab: reads are coalesced along HW, writes
// are coalesced along the (padded) channel pitch.
__global__ void __launch_bounds__(256)
quant_nchw_to_nhwc_kernel(const float* __restrict__ x, int8_t* __restrict__ y, int C, long long HW,
                          int ld, float scale, float inv, int round_mode) {
  extern __shared__ __align__(16) int8_t slab[];  // [64][ld]
  const int n = blockIdx.y;
  const long long hw0 = (long long)blockIdx.x * 64;
  const int npx = (int)((HW - hw0) < 64 ? (HW - hw0) : 64);
  const float* xs = x + (long long)n * C * HW;
  for (int i = threadIdx.x; i < 64 * ld; i += 256) {
    const int c = i >> 6, px = i & 63;  // px fastest: coalesced global reads
    if (c < ld) {
      int8_t q = 0;
      if (c < C && px < npx)
        q = (int8_t)round_sat_i8(quant_pre(__ldg(xs + (long long)c * HW + hw0 + px), scale, inv, round_mode),
                                 round_mode);
      slab[px * ld + c] = q;
    }
  }
  __syncthreads();
  int8_t* yo = y + ((long long)n * HW + hw0) * ld;
  const int bytes = npx * ld;
  if ((ld & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
    for (int i = threadIdx.x; i < (bytes >> 4); i += 256)
      reinterpret_cast<uint4*>(yo)[i] = reinterpret_cast<const uint4*>(slab)[i];
  } else if ((ld & 3) == 0 && (reinterpret_cast<uintptr_t>(y) & 3) == 0) {
    for (int i = threadIdx.x; i < (bytes >> 2); i += 256)
      reinterpret_cast<uint32_t*>(yo)[i] = reinterpret_cast<const uint32_t*>(slab)[i];
  } else {
    for (int i = threadIdx.x; i < bytes; i += 256) yo[i] = slab[i];
  }
}

// ------------------------------------------------------------------ im2col
// Explicit im2col for thin-channel first layers (C <= 4 stored as one 32-bit word
// per pixel): A[m][(r*S+s)*cw + j] = x[n, ih, iw, j], zero outside the image and in
// the [R*S*cw*4, ldA) pad.  One thread produces 16 output bytes (4 words).  The conv
// then runs as a plain tensor-core GEMM with fat 128-byte K rows instead of 49 TMA
// im2col loads of 16 useful bytes each.
__global__ void __launch_bounds__(256)
im2col_words_kernel(const uint32_t* __restrict__ x, uint4* __restrict__ A, int H, int W, int ldxw,
                    int cw, int S, int sh, int sw, int pt, int pl_, int dh, int dw, FastDiv dP, FastDiv dQ,
                    FastDiv dG, uint32_t total, int kwords) {
  // per-word lookup (r*dh, s*dw, c) built once per block
  __shared__ int tab[256];
  for (int kw = threadIdx.x; kw < 256; kw += 256) {
    int v = -1;
    if (kw < kwords) {
      const int tap = kw / cw, c = kw - tap * cw;
      const int r = tap / S, s = tap - r * S;
      v = (r * dh) | ((s * dw) << 10) | (c << 20);
    }
    tab[kw] = v;
  }
  __syncthreads();
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const uint32_t m = fdiv(i, dG);
    const int g = (int)(i - m * dG.d);
    const uint32_t t = fdiv(m, dQ);
    const int oq = (int)(m - t * dQ.d);
    const uint32_t n = fdiv(t, dP);
    const int op = (int)(t - n * dP.d);
    const int ih0 = op * sh - pt, iw0 = oq * sw - pl_;
    const uint32_t* xn = x + (size_t)n * H * W * ldxw;
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = tab[g * 4 + j];
      uint32_t v = 0;
      if (e >= 0) {
        const int ih = ih0 + (e & 1023), iw = iw0 + ((e >> 10) & 1023);
        if ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W)
          v = __ldg(xn + ((size_t)ih * W + iw) * ldxw + (e >> 20));
      }
      o[j] = v;
    }
    A[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// fp32 NCHW -> int8 NHWC with one 32-bit word per pixel (C <= 4): thread = 4 pixels
__global__ void __launch_bounds__(256)
quant_nchw_to_word_kernel(const float* __restrict__ x, uint4* __restrict__ y, int C, long long HW4,
                          float scale, float inv, int round_mode) {
  const int n = blockIdx.y;
  const float* xs = x + (long long)n * C * HW4 * 4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < HW4;
       i += (long long)gridDim.x * blockDim.x) {
    int q[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < C) f = __ldg(reinterpret_cast<const float4*>(xs + (long long)c * HW4 * 4) + i);
      q[0][c] = c < C ? round_sat_i8(quant_pre(f.x, scale, inv, round_mode), round_mode) : 0;
      q[1][c] = c < C ? round_sat_i8(quant_pre(f.y, scale, inv, round_mode), round_mode) : 0;
      q[2][c] = c < C ? round_sat_i8(quant_pre(f.z, scale, inv, round_mode), round_mode) : 0;
      q[3][c] = c < C ? round_sat_i8(quant_pre(f.w, scale, inv, round_mode), round_mode) : 0;
    }
    y[(long long)n * HW4 + i] =
        make_uint4(pack4_i8(q[0][0], q[0][1], q[0][2], q[0][3]), pack4_i8(q[1][0], q[1][1], q[1][2], q[1][3]),
                   pack4_i8(q[2][0], q[2][1], q[2][2], q[2][3]), pack4_i8(q[3][0], q[3][1], q[3][2], q[3][3]));
  }
}

// fp32 / fp16 / uint8 NCHW -> int8 space-to-depth pixels (pl_int8.h: pl_s2d_geometry): thread = TWO adjacent
// 16-byte pixels (4 original columns x 2 rows x 4 channels): one 16 / 8 / 4-byte load per channel and row,
// one 32-byte store.  W % 4 != 0 leaves a last single pixel per row to the half-width path.  Only the
// interior is written: the borders stay zero from the caller's memset.
// uint8 feeds carry the image normalisation of lite/utils/cv/image2tensor.cc (x = (u8 - mean[c]) * norm[c])
// in front of the calib step, so a camera / decoder buffer needs a quarter of the fp32 tensor's PCIe bytes.
struct FeedAffine {
  float mean[4], norm[4];
};
template <typename TIn>
__device__ __forceinline__ void feed_load(const TIn* p, bool four, bool vec4, float4& f);
template <>
__device__ __forceinline__ void feed_load<float>(const float* p, bool four, bool vec4, float4& f) {
  if (four && vec4) {
    f = __ldg(reinterpret_cast<const float4*>(p));
  } else {
    const float2 a = __ldg(reinterpret_cast<const float2*>(p));
    f.x = a.x; f.y = a.y;
    if (four) { const float2 b2 = __ldg(reinterpret_cast<const float2*>(p + 2)); f.z = b2.x; f.w = b2.y; }
  }
}
template <>
__device__ __forceinline__ void feed_load<__half>(const __half* p, bool four, bool vec4, float4& f) {
  if (four && vec4) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
    const float2 b2 = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
    f = make_float4(a.x, a.y, b2.x, b2.y);
  } else {
    const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u));
    f.x = a.x; f.y = a.y;
    if (four) {
      const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(p + 2));
      const float2 b2 = __half22float2(*reinterpret_cast<const __half2*>(&v));
      f.z = b2.x; f.w = b2.y;
    }
  }
}
template <>
__device__ __forceinline__ void feed_load<uint8_t>(const uint8_t* p, bool four, bool vec4, float4& f) {
  if (four && vec4) {
    const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(p));
    f = make_float4((float)(u & 255u), (float)((u >> 8) & 255u), (float)((u >> 16) & 255u), (float)(u >> 24));
  } else {
    const uint32_t u = __ldg(reinterpret_cast<const uint16_t*>(p));
    f.x = (float)(u & 255u); f.y = (float)(u >> 8);
    if (four) {
      const uint32_t v = __ldg(reinterpret_cast<const uint16_t*>(p + 2));
      f.z = (float)(v & 255u); f.w = (float)(v >> 8);
    }
  }
}

template <typename TIn, bool AFFINE>
__global__ void __launch_bounds__(256)
quant_nchw_to_s2d_kernel(const TIn* __restrict__ x, uint4* __restrict__ y, int C, int H, int W, int HP, int WS,
                         int pad_s, float scale, float inv, int round_mode, FeedAffine aff) {
  const int n = blockIdx.y;
  const int Ws = W >> 1, Hs = H >> 1, Wp = (Ws + 1) >> 1;  // pixel pairs per row
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= Hs * Wp) return;
  const int hs = item / Wp, wp = item - hs * Wp;
  const int ws = 2 * wp;
  const bool two = ws + 1 < Ws;
  const TIn* xs = x + (long long)n * C * H * W + (long long)(2 * hs) * W + 2 * ws;
  int q[2][4][4];  // [dy][column 0..3][c]
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < C) {
        feed_load<TIn>(xs + (long long)c * H * W + dy * W, two, (W & 3) == 0, f);
        if (AFFINE) {
          f.x = (f.x - aff.mean[c]) * aff.norm[c];
          f.y = (f.y - aff.mean[c]) * aff.norm[c];
          f.z = (f.z - aff.mean[c]) * aff.norm[c];
          f.w = (f.w - aff.mean[c]) * aff.norm[c];
        }
      }
      q[dy][0][c] = c < C ? round_sat_i8(quant_pre(f.x, scale, inv, round_mode), round_mode) : 0;
      q[dy][1][c] = c < C ? round_sat_i8(quant_pre(f.y, scale, inv, round_mode), round_mode) : 0;
      q[dy][2][c] = c < C ? round_sat_i8(quant_pre(f.z, scale, inv, round_mode), round_mode) : 0;
      q[dy][3][c] = c < C ? round_sat_i8(quant_pre(f.w, scale, inv, round_mode), round_mode) : 0;
    }
  }
  uint4* out = y + ((long long)n * HP + hs + pad_s) * WS + ws + pad_s;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    if (h == 1 && !two) break;
    out[h] = make_uint4(pack4_i8(q[0][2 * h][0], q[0][2 * h][1], q[0][2 * h][2], q[0][2 * h][3]),
                        pack4_i8(q[0][2 * h + 1][0], q[0][2 * h + 1][1], q[0][2 * h + 1][2], q[0][2 * h + 1][3]),
                        pack4_i8(q[1][2 * h][0], q[1][2 * h][1], q[1][2 * h][2], q[1][2 * h][3]),
                        pack4_i8(q[1][2 * h + 1][0], q[1][2 * h + 1][1], q[1][2 * h + 1][2], q[1][2 * h + 1][3]));
  }
}

// int8 NHWC (C <= 4 channels, pixel pitch ldx) -> space-to-depth pixels: the entry of the stride-2 stems when the
// int8 tensor already exists (the reference's op chain calib -> layout -> conv of a kCUDA int8 model; the fused
// quantiser above is the one-launch form of the same thing).  Thread = one 16-byte s2d pixel = 2 x 2 original
// pixels; only the interior is written, the borders stay zero from the caller's one-off memset.
__global__ void __launch_bounds__(256)
nhwc_i8_to_s2d_kernel(const int8_t* __restrict__ x, uint4* __restrict__ y, int C, int ldx, int H, int W, int HP,
                      int WS, int pad_s) {
  const int n = blockIdx.y;
  const int Ws = W >> 1, Hs = H >> 1;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= Hs * Ws) return;
  const int hs = item / Ws, ws = item - hs * Ws;
  const int8_t* p = x + (((long long)n * H + 2 * hs) * W + 2 * ws) * ldx;
  uint32_t w[4];
#pragma unroll
  for (int dy = 0; dy < 2; ++dy)
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
      const int8_t* q = p + ((long long)dy * W + dx) * ldx;
      uint32_t v = 0;
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (c < C) v |= (uint32_t)(uint8_t)__ldg(q + c) << (8 * c);
      w[dy * 2 + dx] = v;
    }
  y[((long long)n * HP + hs + pad_s) * WS + ws + pad_s] = make_uint4(w[0], w[1], w[2], w[3]);
}

extern "C" int pl_nhwc_i8_to_s2d_i8(const int8_t* x, int ldx, int8_t* y_s2d, int N, int C, int H, int W,
                                    const pl_s2d_geometry* g, pl_stream_t stream) {
  if (!x || !y_s2d || !g || !g->supported || N <= 0 || C <= 0 || C > 4 || ldx < C) return PL_ERR_ARG;
  if ((H & 1) || (W & 1) || N > 65535) return PL_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(y_s2d) & 15)) return PL_ERR_ALIGN;
  const long long items = (long long)(H / 2) * (W / 2);
  dim3 grid((unsigned)((items + 255) / 256), (unsigned)N);
  nhwc_i8_to_s2d_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<uint4*>(y_s2d), C, ldx, H, W, g->HP,
                                                                g->WS, g->pad_s);
  ++g_launches;
  return launch_status();
}

// channel pitch change of an int8 NHWC tensor: [pixels][ld_src] (C logical channels) -> [pixels][ld_dst], pad
// channels written as zeros.  One thread per 4-byte word of the destination.  (The kernel classes stage an
// activation whose channel count is not a multiple of 16 at a TMA-legal pitch with this; a 2-D memcpy of 9- or
// 18-byte rows runs at a few GB/s.)
__global__ void __launch_bounds__(256)
repitch_i8_kernel(const int8_t* __restrict__ x, uint32_t* __restrict__ y, long long pixels, int C, int ld_src,
                  int words_dst) {
  const long long total = pixels * words_dst;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long pix = i / words_dst;
    const int c0 = (int)(i - pix * words_dst) * 4;
    const int8_t* p = x + pix * ld_src + c0;
    uint32_t v = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c0 + c < C) v |= (uint32_t)(uint8_t)__ldg(p + c) << (8 * c);
    y[i] = v;
  }
}

extern "C" int pl_repitch_i8_nhwc(const int8_t* x, int ld_src, int8_t* y, int ld_dst, int64_t pixels, int C,
                                  pl_stream_t stream) {
  if (!x || !y || pixels <= 0 || C <= 0 || ld_src < C || ld_dst < C || (ld_dst & 3)) return PL_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(y) & 3) return PL_ERR_ALIGN;
  repitch_i8_kernel<<<grid_for(pixels * (ld_dst / 4), 256), 256, 0, (cudaStream_t)stream>>>(
      x, reinterpret_cast<uint32_t*>(y), pixels, C, ld_src, ld_dst / 4);
  ++g_launches;
  return launch_status();
}

// ------------------------------------------------------------------ pooling
__global__ void __launch_bounds__(256)
pool_kernel(const int8_t* __restrict__ x, int ldx, void* __restrict__ y, int ldy, int N, int H, int W,
            int C, int kh, int kw, int sh, int sw, int pt, int pl_, int P, int Q, int pool_type,
            int exclusive, float rescale, int out_dtype, int round_mode, int vec) {
  const int CG = vec ? (C + 15) / 16 : C;
  const long long total = (long long)N * P * Q * CG;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CG);
    long long t = i / CG;
    const int oq = (int)(t % Q);
    t /= Q;
    const int op = (int)(t % P);
    const int n = (int)(t / P);
    int h0 = op * sh - pt, w0 = oq * sw - pl_;
    const int h1 = h0 + kh < H ? h0 + kh : H, w1 = w0 + kw < W ? w0 + kw : W;
    if (h0 < 0) h0 = 0;
    if (w0 < 0) w0 = 0;
    const int cnt = exclusive ? (h1 - h0) * (w1 - w0) : kh * kw;
    const int c0 = vec ? cg * 16 : cg;
    const int nch = vec ? (C - c0 < 16 ? C - c0 : 16) : 1;
    int sum[16], mx[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { sum[j] = 0; mx[j] = -128; }
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) {
        const int8_t* xp = x + (((long long)n * H + h) * W + w) * ldx + c0;
        if (vec) {
          const uint4 v = __ldg(reinterpret_cast<const uint4*>(xp));
          const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int e = (int)(int8_t)((wd[j >> 2] >> (8 * (j & 3))) & 0xff);
            sum[j] += e;
            mx[j] = e > mx[j] ? e : mx[j];
          }
        } else {
          const int e = (int)__ldg(xp);
          sum[0] += e;
          mx[0] = e > mx[0] ? e : mx[0];
        }
      }
    const long long o = (((long long)n * P + op) * Q + oq) * ldy + c0;
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (pool_type == PL_POOL_MAX)
        v[j] = __fmul_rn((float)mx[j], rescale);
      else
        v[j] = __fmul_rn(__fdiv_rn((float)sum[j], (float)cnt), rescale);
    }
    if (out_dtype == PL_OUT_FP32) {
      float* out = reinterpret_cast<float*>(y) + o;
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (j < nch) out[j] = v[j];
    } else {
      int8_t* out = reinterpret_cast<int8_t*>(y) + o;
      int q[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) q[j] = round_sat_i8(v[j], round_mode);
      if (vec && nch == 16 && (ldy & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
        *reinterpret_cast<uint4*>(out) =
            make_uint4(pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]),
                       pack4_i8(q[8], q[9], q[10], q[11]), pack4_i8(q[12], q[13], q[14], q[15]));
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nch) out[j] = (int8_t)q[j];
      }
    }
  }
}

// word-granular pooling: one thread = 4 channels (one 32-bit word) of one output pixel, so
// a warp reads up to 128 contiguous bytes per window tap.  max runs on sign-extended
// s16x2 pairs (PRMT + VIMNMX.S16x2), avg on zero-extended biased u16x2 pairs (plain adds,
// the bias 128*cnt is removed at the end).  Index math uses launch-time magic division.
__device__ __forceinline__ uint32_t prmt_e(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

// V = words per thread (1, or 4 = one 16-byte load per tap); KF = compile-time window edge
// (3 -> fully unrolled 3x3 with every load issued before the first use, 0 -> runtime loops)
template <int POOL, int V, int KF>
__global__ void __launch_bounds__(256)
pool4_kernel(const uint32_t* __restrict__ x, int ldxw, void* __restrict__ y, int ldy, int H, int W, int C,
             int kh, int kw, int sh, int sw, int pt, int pl_, FastDiv dCw, FastDiv dQ, FastDiv dP,
             uint32_t total, int exclusive, float rescale, int out_dtype, int round_mode, int word_store, int reversed) {
  pdl_launch_dependents();
  pdl_wait();
  // grid-stride over the output items (the host caps the grid at a few resident waves: one short CTA per 256 items
  // spent most of the 3x3 / s2 pool of the ResNet stem on CTA turnover, 64 us for 256 MB)
  for (uint32_t idx_ = blockIdx.x * 256u + threadIdx.x; idx_ < total; idx_ += gridDim.x * 256u) [&] {
  const uint32_t idx = (reversed & 1) ? total - 1u - idx_ : idx_;  // pl_set_tile_order
  const uint64_t pol_ef = kL2EvictFirst;
  uint32_t t = fdiv(idx, dCw);
  const int cw = (int)(idx - t * dCw.d) * V;
  uint32_t u = fdiv(t, dQ);
  const int oq = (int)(t - u * dQ.d);
  const uint32_t n = fdiv(u, dP);
  const int op = (int)(u - n * dP.d);
  const int hs = op * sh - pt, ws = oq * sw - pl_;
  const int h1 = min(hs + kh, H), w1 = min(ws + kw, W);
  const int h0 = max(hs, 0), w0 = max(ws, 0);
  const uint32_t* img = x + (size_t)n * (size_t)(H * W) * (size_t)ldxw + cw;
  // avg: channel pairs (0,1) and (2,3) of each word as 16-bit lane sums; max: a01 holds the packed
  // int8 maxima (one VMAX4 per tap and word instead of two unpacks + two 16-bit maxima)
  uint32_t a01[V], a23[V];
#pragma unroll
  for (int v = 0; v < V; ++v) {
    a01[v] = (POOL == PL_POOL_MAX) ? 0x80808080u : 0u;
    a23[v] = 0u;
  }
  auto load = [&](int h, int w, uint32_t (&d)[V]) {
    const uint32_t* p = img + (uint32_t)((h * W + w) * ldxw);
    if (V == 4) {
      const uint4 q = (reversed & 2) ? ldg_v4_hint(p, pol_ef) : __ldg(reinterpret_cast<const uint4*>(p));  // (PL_L2_HINTS bit 5)
      d[0] = q.x; d[V > 1 ? 1 : 0] = q.y; d[V > 2 ? 2 : 0] = q.z; d[V > 3 ? 3 : 0] = q.w;
    } else {
      d[0] = __ldg(p);
    }
  };
  auto reduce = [&](const uint32_t (&d)[V]) {
#pragma unroll
    for (int v = 0; v < V; ++v) {
      if (POOL == PL_POOL_MAX) {
        a01[v] = __vmaxs4(a01[v], d[v]);
      } else {
        const uint32_t b = d[v] ^ 0x80808080u;  // biased to unsigned: lanes cannot carry into each other
        a01[v] += prmt_e(b, 0u, 0x4140);
        a23[v] += prmt_e(b, 0u, 0x4342);
      }
    }
  };
  if (KF > 0) {
    // clamp the tap coordinates (always a valid address), issue all loads, then drop the
    // taps that fell outside the image
    uint32_t d[KF > 0 ? KF * KF : 1][V];
#pragma unroll
    for (int r = 0; r < KF; ++r)
#pragma unroll
      for (int c = 0; c < KF; ++c) load(min(max(hs + r, 0), H - 1), min(max(ws + c, 0), W - 1), d[r * KF + c]);
#pragma unroll
    for (int r = 0; r < KF; ++r)
#pragma unroll
      for (int c = 0; c < KF; ++c) {
        const bool ok = hs + r >= 0 && hs + r < H && ws + c >= 0 && ws + c < W;
        if (ok) reduce(d[r * KF + c]);
      }
  } else {
    for (int h = h0; h < h1; ++h) {
#pragma unroll 4
      for (int w = w0; w < w1; ++w) {
        uint32_t d[V];
        load(h, w, d);
        reduce(d);
      }
    }
  }
  const int cnt_in = (h1 - h0) * (w1 - w0);
  const int cnt = exclusive ? cnt_in : kh * kw;
  const size_t obase = ((size_t)u * dQ.d + oq) * (size_t)ldy;  // u = n * P + op
  if (POOL == PL_POOL_MAX && V == 4 && out_dtype == PL_OUT_INT8 && rescale == 1.0f && round_mode == PL_ROUND_ARM &&
      word_store && cw * 4 + 16 <= C) {
    int8_t* out = reinterpret_cast<int8_t*>(y) + obase + cw * 4;
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) {  // one 16-byte store (see the per-word path below)
      *reinterpret_cast<uint4*>(out) =
          make_uint4(__vmaxs4(a01[0], 0x81818181u), __vmaxs4(a01[V > 1 ? 1 : 0], 0x81818181u),
                     __vmaxs4(a01[V > 2 ? 2 : 0], 0x81818181u), __vmaxs4(a01[V > 3 ? 3 : 0], 0x81818181u));
      return;
    }
  }
#pragma unroll
  for (int v = 0; v < V; ++v) {
    int e[4];
    if (POOL == PL_POOL_MAX) {
      // same scale in and out (the usual max pool): (float)max * 1.0f rounds to max itself, and the
      // ARM convention clamps at -127
      if (out_dtype == PL_OUT_INT8 && rescale == 1.0f && round_mode == PL_ROUND_ARM && word_store &&
          (cw + v) * 4 + 4 <= C) {
        *reinterpret_cast<uint32_t*>(reinterpret_cast<int8_t*>(y) + obase + (cw + v) * 4) =
            __vmaxs4(a01[v], 0x81818181u);
        continue;
      }
      e[0] = (int)(int8_t)(a01[v] & 0xffu);
      e[1] = (int)(int8_t)((a01[v] >> 8) & 0xffu);
      e[2] = (int)(int8_t)((a01[v] >> 16) & 0xffu);
      e[3] = (int)(int8_t)(a01[v] >> 24);
    } else {
      e[0] = (int)(a01[v] & 0xffffu) - 128 * cnt_in;
      e[1] = (int)(a01[v] >> 16) - 128 * cnt_in;
      e[2] = (int)(a23[v] & 0xffffu) - 128 * cnt_in;
      e[3] = (int)(a23[v] >> 16) - 128 * cnt_in;
    }
    float f[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      f[j] = (POOL == PL_POOL_MAX) ? __fmul_rn((float)e[j], rescale)
                                   : __fmul_rn(__fdiv_rn((float)e[j], (float)cnt), rescale);
    const int c0 = (cw + v) * 4;
    if (c0 >= C) break;
    const size_t o = obase + c0;
    const int nch = C - c0 < 4 ? C - c0 : 4;
    if (out_dtype == PL_OUT_FP32) {
      float* out = reinterpret_cast<float*>(y) + o;
      if (word_store && nch == 4) {
        *reinterpret_cast<float4*>(out) = make_float4(f[0], f[1], f[2], f[3]);
      } else {
        for (int j = 0; j < nch; ++j) out[j] = f[j];
      }
    } else {
      int q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q[j] = round_sat_i8(f[j], round_mode);
      int8_t* out = reinterpret_cast<int8_t*>(y) + o;
      if (word_store) {
        *reinterpret_cast<uint32_t*>(out) = pack4_i8(q[0], q[1], q[2], q[3]);
      } else {
        for (int j = 0; j < nch; ++j) out[j] = (int8_t)q[j];
      }
    }
  }
  }();
}

// 3x3 / stride 2 max pooling, two vertically adjacent outputs per thread (16 bytes of channels each): the two windows
// share one input row, so a thread issues 15 instead of 18 sixteen-byte loads for two outputs (the ResNet stem pool
// moves 256 MB and is bound by its loads).  Same-scale int8 in and out (max of int8 is exact; the ARM convention clamps
// at -127).  Items: (image, output row PAIR, output column, 16-byte channel group).
__global__ void __launch_bounds__(256)
pool3x3s2_max_pair_kernel(const uint32_t* __restrict__ x, int ldxw, int8_t* __restrict__ y, int ldy, int H, int W,
                          int pt, int pl_, int P, FastDiv dG, FastDiv dQ, FastDiv dP2, uint32_t total, int reversed) {
  pdl_launch_dependents();
  pdl_wait();
  for (uint32_t idx_ = blockIdx.x * 256u + threadIdx.x; idx_ < total; idx_ += gridDim.x * 256u) {
    const uint32_t idx = reversed ? total - 1u - idx_ : idx_;
    uint32_t t = fdiv(idx, dG);
    const int cw = (int)(idx - t * dG.d) * 4;
    uint32_t u = fdiv(t, dQ);
    const int oq = (int)(t - u * dQ.d);
    const uint32_t n = fdiv(u, dP2);
    const int op = (int)(u - n * dP2.d) * 2;
    const int hs = op * 2 - pt, ws = oq * 2 - pl_;
    const uint32_t* img = x + (size_t)n * (size_t)(H * W) * (size_t)ldxw + cw;
    uint4 d[5][3];
#pragma unroll
    for (int r = 0; r < 5; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int h = min(max(hs + r, 0), H - 1), w = min(max(ws + c, 0), W - 1);
        d[r][c] = __ldg(reinterpret_cast<const uint4*>(img + (uint32_t)((h * W + w) * ldxw)));
      }
    uint4 a0 = make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u), a1 = a0;
#pragma unroll
    for (int r = 0; r < 5; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const bool ok = hs + r >= 0 && hs + r < H && ws + c >= 0 && ws + c < W;
        if (ok && r <= 2) {
          a0.x = __vmaxs4(a0.x, d[r][c].x); a0.y = __vmaxs4(a0.y, d[r][c].y);
          a0.z = __vmaxs4(a0.z, d[r][c].z); a0.w = __vmaxs4(a0.w, d[r][c].w);
        }
        if (ok && r >= 2) {
          a1.x = __vmaxs4(a1.x, d[r][c].x); a1.y = __vmaxs4(a1.y, d[r][c].y);
          a1.z = __vmaxs4(a1.z, d[r][c].z); a1.w = __vmaxs4(a1.w, d[r][c].w);
        }
      }
    const uint32_t lo = 0x81818181u;  // clamp at -127
    const size_t o0 = ((size_t)(n * (uint32_t)P + (uint32_t)op) * dQ.d + (uint32_t)oq) * (size_t)ldy + (size_t)cw * 4;
    *reinterpret_cast<uint4*>(y + o0) =
        make_uint4(__vmaxs4(a0.x, lo), __vmaxs4(a0.y, lo), __vmaxs4(a0.z, lo), __vmaxs4(a0.w, lo));
    if (op + 1 < P)
      *reinterpret_cast<uint4*>(y + o0 + (size_t)dQ.d * (size_t)ldy) =
          make_uint4(__vmaxs4(a1.x, lo), __vmaxs4(a1.y, lo), __vmaxs4(a1.z, lo), __vmaxs4(a1.w, lo));
  }
}

template <int POOL>
static void launch_pool4(int V, int KF, unsigned grid, cudaStream_t st, const uint32_t* x, int ldxw, void* y, int ldy,
                         int H, int W, int C, int kh, int kw, int sh, int sw, int pt, int pl_, FastDiv dCw, FastDiv dQ,
                         FastDiv dP, uint32_t total, int exclusive, float rescale, int out_dtype, int round_mode,
                         int word_store) {
  static const int hints = getenv("PL_L2_HINTS") ? atoi(getenv("PL_L2_HINTS")) : 15;
  const int reversed = (next_tile_order_reversed() ? 1 : 0) | ((hints & 32) ? 2 : 0);  // bit 1: L2 evict_first loads
#define PL_POOL_ARGS x, ldxw, y, ldy, H, W, C, kh, kw, sh, sw, pt, pl_, dCw, dQ, dP, total, exclusive, rescale, \
                     out_dtype, round_mode, word_store, reversed
  if (V == 4 && KF == 3) launch_pdl(pool4_kernel<POOL, 4, 3>, dim3(grid), dim3(256), 0, st, PL_POOL_ARGS);
  else if (V == 4) launch_pdl(pool4_kernel<POOL, 4, 0>, dim3(grid), dim3(256), 0, st, PL_POOL_ARGS);
  else if (KF == 3) launch_pdl(pool4_kernel<POOL, 1, 3>, dim3(grid), dim3(256), 0, st, PL_POOL_ARGS);
  else if (KF == 7) launch_pdl(pool4_kernel<POOL, 1, 7>, dim3(grid), dim3(256), 0, st, PL_POOL_ARGS);
  else launch_pdl(pool4_kernel<POOL, 1, 0>, dim3(grid), dim3(256), 0, st, PL_POOL_ARGS);
#undef PL_POOL_ARGS
}

// ------------------------------------------------------------- residual add
__device__ __forceinline__ float add_value(int a, int b, float sa, float sb, int act, float arg) {
  const float v = __fmaf_rn((float)a, sa, __fmul_rn((float)b, sb));
  return apply_act(v, act, arg);
}

__global__ void __launch_bounds__(256)
add_kernel(const int8_t* __restrict__ a, const int8_t* __restrict__ b, void* __restrict__ y,
           long long n, float sa, float sb, int act, float arg, int out_dtype, int round_mode,
           int vec) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const long long n16 = n >> 4;
    for (; i < n16; i += stride) {
      const uint4 va = __ldg(reinterpret_cast<const uint4*>(a) + i);
      const uint4 vb = __ldg(reinterpret_cast<const uint4*>(b) + i);
      const uint32_t wa[4] = {va.x, va.y, va.z, va.w}, wb[4] = {vb.x, vb.y, vb.z, vb.w};
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int ea = (int)(int8_t)((wa[j >> 2] >> (8 * (j & 3))) & 0xff);
        const int eb = (int)(int8_t)((wb[j >> 2] >> (8 * (j & 3))) & 0xff);
        v[j] = add_value(ea, eb, sa, sb, act, arg);
      }
      if (out_dtype == PL_OUT_FP32) {
        float4* yp = reinterpret_cast<float4*>(y) + i * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) yp[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
      } else {
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          o[k] = pack4_i8(round_sat_i8(v[4 * k], round_mode), round_sat_i8(v[4 * k + 1], round_mode),
                          round_sat_i8(v[4 * k + 2], round_mode), round_sat_i8(v[4 * k + 3], round_mode));
        reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
    i = (n16 << 4) + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  }
  for (; i < n; i += stride) {
    const float v = add_value((int)a[i], (int)b[i], sa, sb, act, arg);
    if (out_dtype == PL_OUT_FP32)
      reinterpret_cast<float*>(y)[i] = v;
    else
      reinterpret_cast<int8_t*>(y)[i] = (int8_t)round_sat_i8(v, round_mode);
  }
}

}  // namespace pl

using namespace pl;

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int pl_quantize_f32_i8(const float* x, int8_t* y, int64_t n, float scale, int round_mode,
                                  pl_stream_t stream) {
  if (!x || !y || n < 0 || !(scale > 0.f)) return PL_ERR_ARG;
  if (n == 0) return PL_OK;
  const int vec = al16(x) && al16(y);
  quantize_kernel<<<grid_for(vec ? (n + 15) / 16 : n, 256), 256, 0, (cudaStream_t)stream>>>(
      x, y, n, scale, 1.f / scale, round_mode, vec);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_dequantize_i8_f32(const int8_t* x, float* y, int64_t n, float scale,
                                    pl_stream_t stream) {
  if (!x || !y || n < 0) return PL_ERR_ARG;
  if (n == 0) return PL_OK;
  const int vec = al16(x) && al16(y);
  dequantize_kernel<<<grid_for(vec ? (n + 15) / 16 : n, 256), 256, 0, (cudaStream_t)stream>>>(
      x, y, n, scale, vec);
  ++g_launches;
  return launch_status();
}

template <typename T>
static int launch_transpose(const T* src, T* dst, int B, int rows, int cols, long long ld_src,
                            long long ld_dst, long long bs_src, long long bs_dst, cudaStream_t st) {
  if (B > 65535) return PL_ERR_UNSUPPORTED;
  dim3 grid((cols + 63) / 64, (rows + 63) / 64, B);
  transpose_kernel<T><<<grid, 256, 0, st>>>(src, dst, rows, cols, ld_src, ld_dst, bs_src, bs_dst);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_transpose_i8(const int8_t* src, int ld_src, int8_t* dst, int ld_dst, int rows,
                               int cols, pl_stream_t stream) {
  if (!src || !dst || rows <= 0 || cols <= 0 || ld_src < cols || ld_dst < rows) return PL_ERR_ARG;
  return launch_transpose<int8_t>(src, dst, 1, rows, cols, ld_src, ld_dst, 0, 0, (cudaStream_t)stream);
}

// int8 [N][C][HW] -> [N][HW][ld] for C <= 4 (the image feed of a quantised net after its calib): a 64 x 64 tile
// transpose would use 3 of its 64 rows.  Thread = 4 consecutive pixels: one 32-bit load per channel plane, the
// 4 * ld output bytes assembled in registers (pad channels zero) and stored as words when the pitch allows.
__global__ void __launch_bounds__(256)
nchw_to_nhwc_thin_i8_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ y, int C, long long HW, int ld) {
  const int n = blockIdx.y;
  const long long q4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // group of 4 pixels
  if (q4 * 4 >= HW) return;
  const int8_t* xs = x + (long long)n * C * HW + q4 * 4;
  uint32_t plane[4] = {0, 0, 0, 0};
#pragma unroll
  for (int c = 0; c < 4; ++c)
    if (c < C) plane[c] = __ldg(reinterpret_cast<const uint32_t*>(xs + (long long)c * HW));
  int8_t* out = y + ((long long)n * HW + q4 * 4) * ld;
  uint8_t b[16];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int c = 0; c < 4; ++c) b[p * 4 + c] = (uint8_t)(plane[c] >> (8 * p));
  if (ld == 4) {
    *reinterpret_cast<uint4*>(out) = *reinterpret_cast<const uint4*>(b);
  } else if (ld == 3) {  // 12 contiguous bytes, 4-byte aligned (4 pixels x 3)
    const uint32_t w0 = b[0] | (b[1] << 8) | (b[2] << 16) | ((uint32_t)b[4] << 24);
    const uint32_t w1 = b[5] | (b[6] << 8) | (b[8] << 16) | ((uint32_t)b[9] << 24);
    const uint32_t w2 = b[10] | (b[12] << 8) | (b[13] << 16) | ((uint32_t)b[14] << 24);
    uint32_t* o = reinterpret_cast<uint32_t*>(out);
    o[0] = w0; o[1] = w1; o[2] = w2;
  } else {
#pragma unroll
    for (int p = 0; p < 4; ++p)
      for (int c = 0; c < ld; ++c) out[p * ld + c] = c < 4 ? (int8_t)b[p * 4 + c] : (int8_t)0;
  }
}

extern "C" int pl_nchw_to_nhwc(const void* x, void* y, int N, int C, int H, int W, int ld,
                               int elem_bytes, pl_stream_t stream) {
  if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || ld < C) return PL_ERR_ARG;
  const long long HW = (long long)H * W;
  if (HW > 0x7fffffffLL) return PL_ERR_UNSUPPORTED;
  if (elem_bytes == 1 && C <= 4 && (HW % 4) == 0 && N <= 65535 && (reinterpret_cast<uintptr_t>(x) & 3) == 0 &&
      (reinterpret_cast<uintptr_t>(y) & 15) == 0 && (ld == C || ld == 4)) {
    dim3 grid((unsigned)((HW / 4 + 255) / 256), (unsigned)N);
    nchw_to_nhwc_thin_i8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const int8_t*)x, (int8_t*)y, C, HW, ld);
    ++g_launches;
    return launch_status();
  }
  int rc;
  // [N][C][HW] -> [N][HW][ld]
  if (elem_bytes == 1)
    rc = launch_transpose<int8_t>((const int8_t*)x, (int8_t*)y, N, C, (int)HW, HW, ld, (long long)C * HW,
                                  HW * ld, (cudaStream_t)stream);
  else if (elem_bytes == 4)
    rc = launch_transpose<float>((const float*)x, (float*)y, N, C, (int)HW, HW, ld, (long long)C * HW,
                                 HW * ld, (cudaStream_t)stream);
  else
    return PL_ERR_ARG;
  if (rc == PL_OK && ld > C) {
    zero_pad_kernel<<<grid_for((long long)N * HW * (ld - C) * elem_bytes, 256), 256, 0,
                      (cudaStream_t)stream>>>((int8_t*)y, (long long)N * HW, C, ld, elem_bytes);
    ++g_launches;
    rc = launch_status();
  }
  return rc;
}

extern "C" int pl_nhwc_to_nchw(const void* x, void* y, int N, int C, int H, int W, int ld,
                               int elem_bytes, pl_stream_t stream) {
  if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || ld < C) return PL_ERR_ARG;
  const long long HW = (long long)H * W;
  if (HW > 0x7fffffffLL) return PL_ERR_UNSUPPORTED;
  // [N][HW][ld] -> [N][C][HW]
  if (elem_bytes == 1)
    return launch_transpose<int8_t>((const int8_t*)x, (int8_t*)y, N, (int)HW, C, ld, HW, HW * ld,
                                    (long long)C * HW, (cudaStream_t)stream);
  if (elem_bytes == 4)
    return launch_transpose<float>((const float*)x, (float*)y, N, (int)HW, C, ld, HW, HW * ld,
                                   (long long)C * HW, (cudaStream_t)stream);
  return PL_ERR_ARG;
}

extern "C" int pl_quantize_nchw_f32_to_nhwc_i8(const float* x, int8_t* y, int N, int C, int H, int W,
                                               int ldy, float scale, int round_mode,
                                               pl_stream_t stream) {
  if (!x || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || ldy < C || !(scale > 0.f)) return PL_ERR_ARG;
  if (N > 65535 || (size_t)64 * ldy > 48 * 1024) return PL_ERR_UNSUPPORTED;
  const long long HW = (long long)H * W;
  if (ldy == 4 && C <= 4 && (HW % 4) == 0 && al16(x) && al16(y)) {
    dim3 grid((unsigned)((HW / 4 + 255) / 256 > 64 ? 64 : (HW / 4 + 255) / 256), (unsigned)N);
    quant_nchw_to_word_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<uint4*>(y), C, HW / 4,
                                                                      scale, 1.f / scale, round_mode);
    ++g_launches;
    return launch_status();
  }
  dim3 grid((unsigned)((HW + 63) / 64), (unsigned)N);
  quant_nchw_to_nhwc_kernel<<<grid, 256, (size_t)64 * ldy, (cudaStream_t)stream>>>(
      x, y, C, HW, ldy, scale, 1.f / scale, round_mode);
  ++g_launches;
  return launch_status();
}

template <typename TIn, bool AFFINE>
static int launch_quant_s2d(const TIn* x, int8_t* y_s2d, int N, int C, int H, int W, const pl_s2d_geometry* g,
                            float scale, int round_mode, const FeedAffine& aff, pl_stream_t stream) {
  if (!x || !y_s2d || !g || !g->supported || N <= 0 || C <= 0 || C > 4 || !(scale > 0.f)) return PL_ERR_ARG;
  if ((H & 1) || (W & 1) || N > 65535) return PL_ERR_UNSUPPORTED;
  // rows are read with 4-element vector loads when W % 4 == 0: the base must be aligned to that vector
  if (!al16(y_s2d) || (reinterpret_cast<uintptr_t>(x) & (4 * sizeof(TIn) - 1))) return PL_ERR_ALIGN;
  const long long items = (long long)(H / 2) * ((W / 2 + 1) / 2);
  dim3 grid((unsigned)((items + 255) / 256), (unsigned)N);
  quant_nchw_to_s2d_kernel<TIn, AFFINE><<<grid, 256, 0, (cudaStream_t)stream>>>(
      x, reinterpret_cast<uint4*>(y_s2d), C, H, W, g->HP, g->WS, g->pad_s, scale, 1.f / scale, round_mode, aff);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_quantize_nchw_f32_to_s2d_i8(const float* x, int8_t* y_s2d, int N, int C, int H, int W,
                                              const pl_s2d_geometry* g, float scale, int round_mode,
                                              pl_stream_t stream) {
  return launch_quant_s2d<float, false>(x, y_s2d, N, C, H, W, g, scale, round_mode, FeedAffine{}, stream);
}

extern "C" int pl_quantize_nchw_f16_to_s2d_i8(const void* x_f16, int8_t* y_s2d, int N, int C, int H, int W,
                                              const pl_s2d_geometry* g, float scale, int round_mode,
                                              pl_stream_t stream) {
  return launch_quant_s2d<__half, false>(reinterpret_cast<const __half*>(x_f16), y_s2d, N, C, H, W, g, scale,
                                         round_mode, FeedAffine{}, stream);
}

extern "C" int pl_quantize_nchw_u8_to_s2d_i8(const uint8_t* x_u8, int8_t* y_s2d, int N, int C, int H, int W,
                                             const pl_s2d_geometry* g, const float* mean, const float* norm,
                                             float scale, int round_mode, pl_stream_t stream) {
  if (!mean || !norm || C <= 0 || C > 4) return PL_ERR_ARG;
  FeedAffine aff{};
  for (int c = 0; c < C; ++c) {
    aff.mean[c] = mean[c];
    aff.norm[c] = norm[c];
  }
  return launch_quant_s2d<uint8_t, true>(x_u8, y_s2d, N, C, H, W, g, scale, round_mode, aff, stream);
}

// adaptive pooling (PoolParam::adaptive, lite/kernels/cuda/pool_compute.cu:69-187): one thread per output
// element, window [floor(i * in / out), ceil((i + 1) * in / out)); consecutive threads take consecutive channels
__global__ void __launch_bounds__(256)
adaptive_pool_kernel(const int8_t* __restrict__ x, int ldx, void* __restrict__ y, int ldy, int N, int H, int W, int C,
                     int P, int Q, int pool_type, float rescale, int out_dtype, int round_mode) {
  const long long total = (long long)N * P * Q * C;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    long long t = i / C;
    const int q = (int)(t % Q);
    t /= Q;
    const int p = (int)(t % P);
    const int n = (int)(t / P);
    const int h0 = (p * H) / P, h1 = ((p + 1) * H + P - 1) / P;
    const int w0 = (q * W) / Q, w1 = ((q + 1) * W + Q - 1) / Q;
    int sum = 0, mx = -128;
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) {
        const int v = __ldg(x + (((long long)n * H + h) * W + w) * ldx + c);
        sum += v;
        mx = v > mx ? v : mx;
      }
    const float cnt = (float)((h1 - h0) * (w1 - w0));
    const float v = pool_type == PL_POOL_MAX ? (float)mx * rescale : __fdiv_rn((float)sum, cnt) * rescale;
    const long long o = (((long long)n * P + p) * Q + q) * ldy + c;
    if (out_dtype == PL_OUT_FP32)
      reinterpret_cast<float*>(y)[o] = v;
    else
      reinterpret_cast<int8_t*>(y)[o] = (int8_t)round_sat_i8(v, round_mode);
  }
}

extern "C" int pl_adaptive_pool2d_i8_nhwc(const int8_t* x, int ldx, void* y, int ldy, int N, int H, int W, int C,
                                          int P, int Q, int pool_type, float rescale, int out_dtype, int round_mode,
                                          pl_stream_t stream) {
  if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0 || P <= 0 || Q <= 0 || ldx < C || ldy < C) return PL_ERR_ARG;
  if (out_dtype != PL_OUT_INT8 && out_dtype != PL_OUT_FP32) return PL_ERR_ARG;
  if (pool_type != PL_POOL_MAX && pool_type != PL_POOL_AVG) return PL_ERR_ARG;
  const long long total = (long long)N * P * Q * C;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  adaptive_pool_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, y, ldy, N, H, W, C, P, Q, pool_type,
                                                                          rescale, out_dtype, round_mode);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_pool2d_i8_nhwc(const int8_t* x, int ldx, void* y, int ldy, int N, int H, int W, int C,
                                 int kh, int kw, int sh, int sw, int pt, int pl_, int P, int Q,
                                 int pool_type, int exclusive, float rescale, int out_dtype,
                                 int round_mode, pl_stream_t stream) {
  if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0 || kh <= 0 || kw <= 0 || sh <= 0 || sw <= 0 ||
      P <= 0 || Q <= 0 || ldx < C || ldy < C)
    return PL_ERR_ARG;
  if (out_dtype != PL_OUT_INT8 && out_dtype != PL_OUT_FP32) return PL_ERR_ARG;
  // word-granular fast path: 4-byte aligned input pixels, window small enough for the
  // 16-bit lane sums, 32-bit item index
  const int Cw = (C + 3) / 4;
  if ((reinterpret_cast<uintptr_t>(x) & 3) == 0 && ldx % 4 == 0 && ldx >= Cw * 4 && kh * kw <= 255 &&
      (long long)N * P * Q * Cw < (1ll << 31) && (long long)H * W * (ldx / 4) < (1ll << 31)) {
    // 16-byte taps when every pixel is 16-byte aligned and covers whole 4-word groups
    // (only when that still leaves a few waves of threads: global pooling stays word-wide)
    const int V = (al16(x) && ldx % 16 == 0 && ldx >= (C + 15) / 16 * 16 && C >= 16 &&
                   (long long)N * P * Q * ((Cw + 3) / 4) >= 148ll * 2048 * 2)
                      ? 4
                      : 1;
    // 7x7 (the global pools of ResNet / MobileNet at 224): word-wide, all 49 taps in flight per thread
    // (the runtime loop serialised ~49 memory latencies: 17 us for 26 MB)
    const int KF = (kh == 3 && kw == 3) ? 3 : ((kh == 7 && kw == 7 && V == 1) ? 7 : 0);
    const int groups = (Cw + V - 1) / V;
    const long long witems = (long long)N * P * Q * groups;
    // whole-word stores need the pad channels of the last word to be addressable
    const int word_store = (ldy >= Cw * 4) && (ldy % 4 == 0) &&
                           ((reinterpret_cast<uintptr_t>(y) & (out_dtype == PL_OUT_INT8 ? 3 : 15)) == 0);
    unsigned grid = (unsigned)((witems + 255) / 256);
    {
      // at most PL_POOL_WAVES resident waves of CTAs (8 CTAs of 256 threads per SM); the kernel strides over the rest
      int sms = 148, dev = 0;
      cudaGetDevice(&dev);
      static int sms_of[64] = {};
      if (sms_of[dev & 63] == 0) cudaDeviceGetAttribute(&sms_of[dev & 63], cudaDevAttrMultiProcessorCount, dev);
      if (sms_of[dev & 63] > 0) sms = sms_of[dev & 63];
      static const int waves = getenv("PL_POOL_WAVES") ? atoi(getenv("PL_POOL_WAVES")) : 2;
      const unsigned cap = (unsigned)(sms * 8 * (waves > 0 ? waves : 1));
      if (waves > 0 && grid > cap) grid = cap;
    }
    const FastDiv dCw = make_fastdiv((uint32_t)groups), dQ = make_fastdiv((uint32_t)Q), dP = make_fastdiv((uint32_t)P);
    {
      // (ResNet stem pool at b256: 62.4 -> 56.7 us; PL_POOL_PAIR=0 keeps one output per thread)
      static const int pair_knob = getenv("PL_POOL_PAIR") ? atoi(getenv("PL_POOL_PAIR")) : 1;
      if (pair_knob && pool_type == PL_POOL_MAX && V == 4 && KF == 3 && sh == 2 && sw == 2 && out_dtype == PL_OUT_INT8 &&
          rescale == 1.0f && round_mode == PL_ROUND_ARM && C % 16 == 0 && ldy % 16 == 0 &&
          (reinterpret_cast<uintptr_t>(y) & 15) == 0 && pt >= 0 && pt <= 2 && pl_ >= 0 && pl_ <= 2) {
        const int P2 = (P + 1) / 2;
        const long long items = (long long)N * P2 * Q * groups;
        unsigned g2 = (unsigned)((items + 255) / 256);
        if (g2 > grid) g2 = grid;  // (grid is already capped at the resident waves)
        launch_pdl(pool3x3s2_max_pair_kernel, dim3(g2), dim3(256), 0, (cudaStream_t)stream,
                   reinterpret_cast<const uint32_t*>(x), ldx / 4, reinterpret_cast<int8_t*>(y), ldy, H, W, pt, pl_, P, dCw, dQ,
                   make_fastdiv((uint32_t)P2), (uint32_t)items, next_tile_order_reversed() ? 1 : 0);
        ++g_launches;
        return launch_status();
      }
    }
    if (pool_type == PL_POOL_MAX)
      launch_pool4<PL_POOL_MAX>(V, KF, grid, (cudaStream_t)stream, reinterpret_cast<const uint32_t*>(x), ldx / 4, y,
                                ldy, H, W, C, kh, kw, sh, sw, pt, pl_, dCw, dQ, dP, (uint32_t)witems, exclusive,
                                rescale, out_dtype, round_mode, word_store);
    else
      launch_pool4<PL_POOL_AVG>(V, KF, grid, (cudaStream_t)stream, reinterpret_cast<const uint32_t*>(x), ldx / 4, y,
                                ldy, H, W, C, kh, kw, sh, sw, pt, pl_, dCw, dQ, dP, (uint32_t)witems, exclusive,
                                rescale, out_dtype, round_mode, word_store);
    ++g_launches;
    return launch_status();
  }
  const int vec = al16(x) && (ldx % 16 == 0);
  const long long items = (long long)N * P * Q * (vec ? (C + 15) / 16 : C);
  pool_kernel<<<grid_for(items, 256), 256, 0, (cudaStream_t)stream>>>(
      x, ldx, y, ldy, N, H, W, C, kh, kw, sh, sw, pt, pl_, P, Q, pool_type, exclusive, rescale,
      out_dtype, round_mode, vec);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_add_i8(const int8_t* a, const int8_t* b, void* y, int64_t n, float sa, float sb,
                         int act, float act_arg, int out_dtype, int round_mode, pl_stream_t stream) {
  if (!a || !b || !y || n < 0) return PL_ERR_ARG;
  if (out_dtype != PL_OUT_INT8 && out_dtype != PL_OUT_FP32) return PL_ERR_ARG;
  if (n == 0) return PL_OK;
  const int vec = al16(a) && al16(b) && al16(y);
  add_kernel<<<grid_for(vec ? (n + 15) / 16 : n, 256), 256, 0, (cudaStream_t)stream>>>(
      a, b, y, n, sa, sb, act, act_arg, out_dtype, round_mode, vec);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_im2col_i8_nhwc(const int8_t* x, int ldx, int8_t* A, int ldA, int N, int H, int W, int cbytes,
                                 int R, int S, int sh, int sw, int pt, int pl_, int dh, int dw, int P, int Q,
                                 pl_stream_t stream) {
  if (!x || !A || N <= 0 || H <= 0 || W <= 0 || cbytes <= 0 || R <= 0 || S <= 0 || sh <= 0 || sw <= 0 ||
      dh <= 0 || dw <= 0 || P <= 0 || Q <= 0)
    return PL_ERR_ARG;
  if ((cbytes % 4) || (ldx % 4) || ldx < cbytes || (ldA % 16) || ldA < R * S * cbytes) return PL_ERR_ALIGN;
  if ((reinterpret_cast<uintptr_t>(x) & 3) || !al16(A)) return PL_ERR_ALIGN;
  const long long M = (long long)N * P * Q;
  const int groups = ldA / 16;
  const int kwords = R * S * (cbytes / 4);
  if (M * groups >= (1ll << 31) || kwords > 256 || (R - 1) * dh > 1023 || (S - 1) * dw > 1023)
    return PL_ERR_UNSUPPORTED;
  im2col_words_kernel<<<grid_for(M * groups, 256, 16), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint32_t*>(x), reinterpret_cast<uint4*>(A), H, W, ldx / 4, cbytes / 4, S, sh, sw,
      pt, pl_, dh, dw, make_fastdiv((uint32_t)P), make_fastdiv((uint32_t)Q), make_fastdiv((uint32_t)groups),
      (uint32_t)(M * groups), kwords);
  ++g_launches;
  return launch_status();
}
