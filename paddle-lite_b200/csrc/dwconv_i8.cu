// dwconv_i8.cu -- depthwise int8 conv over NHWC (groups == C == K).
//
// Bandwidth-bound (9 MAC per 2 bytes): no tensor cores.  The reference has no
// kCUDA int8 depthwise at all (cuDNN int8 path ignores groups,
// lite/backends/cuda/math/cudnn_conv.cc:320-585); semantics follow the ARM
// kernels lite/backends/arm/math/conv3x3s{1,2}_depthwise_int8.cc /
// conv5x5s{1,2}_depthwise_int8.cc: exact int32 accumulation, then the common
// epilogue (common.cuh).  Filter layout as the reference stores it: [C][R][S].
//
// Fast path (3x3, stride 1 or 2, dilation 1): one thread owns 4 channels (one
// 32-bit word per pixel, so a warp reads 128 contiguous bytes) and two adjacent
// output columns.  The 4 (s1) or 5 (s2) input words of a filter row are byte
// transposed with 8 PRMTs into one word per channel holding 4 consecutive
// columns; a 3-tap row of the filter is then ONE dp4a against a pre-shifted
// weight word.  Filter words, scale and bias live in registers for the whole
// grid-stride loop.
// Generic path (any R,S,stride,dilation): one dp4a per MAC against a byte-masked
// weight word.
#include "common.cuh"

namespace pl {

struct DwParams {
  const int8_t* x;
  const int8_t* w;  // [C][R][S]
  void* y;
  int N, H, W, C, ldx, ldy, P, Q;
  int R, S, sh, sw, pt, pl_, dh, dw;
  int CG;       // channel groups of 4
  long long npairs;  // N * P * ceil(Q/2)   (fast path)
  long long npix;    // N * P * Q           (generic path)
  EpiParams epi;
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

__device__ __forceinline__ uint32_t ldg_u32(const int8_t* p) {
  return __ldg(reinterpret_cast<const unsigned int*>(p));
}

// write 4 channels of one output pixel
__device__ __forceinline__ void dw_store4(const DwParams& p, const int (&acc)[4], const float (&sc)[4],
                                          const float (&bs)[4], long long pix, int c0) {
  const EpiParams& e = p.epi;
  const int nvalid = p.C - c0 < 4 ? p.C - c0 : 4;
  if (e.out_dtype == PL_OUT_INT32) {
    int* out = reinterpret_cast<int*>(p.y) + pix * p.ldy + c0;
    for (int j = 0; j < nvalid; ++j) out[j] = acc[j];
    return;
  }
  float v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float t = __fmaf_rn(small_int2float(acc[j]), sc[j], bs[j]);
    v[j] = (e.round_mode == PL_ROUND_CUDA_LEGACY) ? apply_act_legacy(t, e.act, e.act_arg, e.vec4)
                                                  : apply_act(t, e.act, e.act_arg);
  }
  if (e.out_dtype == PL_OUT_FP32) {
    float* out = reinterpret_cast<float*>(p.y) + pix * p.ldy + c0;
    if (nvalid == 4 && (p.ldy & 3) == 0)
      *reinterpret_cast<float4*>(out) = make_float4(v[0], v[1], v[2], v[3]);
    else
      for (int j = 0; j < nvalid; ++j) out[j] = v[j];
  } else {
    int q[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) q[j] = round_sat_i8(v[j], e.round_mode);
    int8_t* out = reinterpret_cast<int8_t*>(p.y) + pix * p.ldy + c0;
    if (nvalid == 4 && (p.ldy & 3) == 0)
      *reinterpret_cast<uint32_t*>(out) = pack4_i8(q[0], q[1], q[2], q[3]);
    else
      for (int j = 0; j < nvalid; ++j) out[j] = (int8_t)q[j];
  }
}

// ---------------------------------------------------------------------------
// fast path: 3x3, stride STRIDE in {1,2}, dilation 1
// block = (CGB channel groups) x (256/CGB pixel pairs); grid.y = channel-group blocks
// ---------------------------------------------------------------------------
template <int STRIDE>
__global__ void __launch_bounds__(256) dw3x3_kernel(const DwParams p, int CGB) {
  const int cg = blockIdx.y * CGB + (threadIdx.x % CGB);
  const int pslot = threadIdx.x / CGB;
  const int pper = 256 / CGB;
  if (cg >= p.CG) return;
  const int c0 = cg * 4;

  // filter words: wa[r][c] = {w0,w1,w2,0}; wb = taps shifted for the 2nd column
  // (s1: {0,w0,w1,w2}; s2: {0,0,w0,w1} plus w2 against the 5th word).
  uint32_t wa[3][4], wb[3][4], wc[3][4];
  float sc[4], bs[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + j;
    const bool ok = c < p.C;
    sc[j] = (ok && p.epi.scale) ? __ldg(p.epi.scale + c) : 0.f;
    bs[j] = (ok && p.epi.bias) ? __ldg(p.epi.bias + c) : 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      uint32_t w0 = 0, w1 = 0, w2 = 0;
      if (ok) {
        const int8_t* wp = p.w + (long long)c * 9 + r * 3;
        w0 = (uint8_t)__ldg(wp);
        w1 = (uint8_t)__ldg(wp + 1);
        w2 = (uint8_t)__ldg(wp + 2);
      }
      wa[r][j] = w0 | (w1 << 8) | (w2 << 16);
      if (STRIDE == 1) {
        wb[r][j] = (w0 << 8) | (w1 << 16) | (w2 << 24);
        wc[r][j] = 0;
      } else {
        wb[r][j] = (w0 << 16) | (w1 << 24);
        wc[r][j] = w2 << (8 * j);  // masked: picks channel j of the raw 5th word
      }
    }
  }

  const int Q2 = (p.Q + 1) >> 1;
  for (long long pp = (long long)blockIdx.x * pper + pslot; pp < p.npairs;
       pp += (long long)gridDim.x * pper) {
    const int q0 = (int)(pp % Q2) * 2;
    const long long t = pp / Q2;
    const int op = (int)(t % p.P);
    const int n = (int)(t / p.P);
    const int iw0 = q0 * STRIDE - p.pl_;
    const int ih0 = op * STRIDE - p.pt;
    int acc0[4] = {0, 0, 0, 0}, acc1[4] = {0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = ih0 + r;
      if (ih < 0 || ih >= p.H) continue;
      const int8_t* row = p.x + ((long long)n * p.H + ih) * p.W * p.ldx + c0;
      uint32_t xw[5];
#pragma unroll
      for (int i = 0; i < (STRIDE == 1 ? 4 : 5); ++i) {
        const int iw = iw0 + i;
        xw[i] = (iw >= 0 && iw < p.W) ? ldg_u32(row + (long long)iw * p.ldx) : 0u;
      }
      // 4x4 byte transpose: t[c] = {x0.c, x1.c, x2.c, x3.c}
      const uint32_t a = prmt(xw[0], xw[1], 0x5140), b = prmt(xw[2], xw[3], 0x5140);
      const uint32_t c_ = prmt(xw[0], xw[1], 0x7362), d = prmt(xw[2], xw[3], 0x7362);
      uint32_t tr[4];
      tr[0] = prmt(a, b, 0x5410);
      tr[1] = prmt(a, b, 0x7632);
      tr[2] = prmt(c_, d, 0x5410);
      tr[3] = prmt(c_, d, 0x7632);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc0[j] = __dp4a((int)tr[j], (int)wa[r][j], acc0[j]);
        acc1[j] = __dp4a((int)tr[j], (int)wb[r][j], acc1[j]);
        if (STRIDE == 2) acc1[j] = __dp4a((int)xw[4], (int)wc[r][j], acc1[j]);
      }
    }
    const long long pix = ((long long)n * p.P + op) * p.Q + q0;
    dw_store4(p, acc0, sc, bs, pix, c0);
    if (q0 + 1 < p.Q) dw_store4(p, acc1, sc, bs, pix + 1, c0);
  }
}

// ---------------------------------------------------------------------------
// generic path
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dw_generic_kernel(const DwParams p, int CGB, int aligned) {
  const int cg = blockIdx.y * CGB + (threadIdx.x % CGB);
  const int pslot = threadIdx.x / CGB;
  const int pper = 256 / CGB;
  if (cg >= p.CG) return;
  const int c0 = cg * 4;
  const int RS = p.R * p.S;
  float sc[4], bs[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + j;
    const bool ok = c < p.C;
    sc[j] = (ok && p.epi.scale) ? __ldg(p.epi.scale + c) : 0.f;
    bs[j] = (ok && p.epi.bias) ? __ldg(p.epi.bias + c) : 0.f;
  }
  for (long long px = (long long)blockIdx.x * pper + pslot; px < p.npix;
       px += (long long)gridDim.x * pper) {
    const int oq = (int)(px % p.Q);
    const long long t = px / p.Q;
    const int op = (int)(t % p.P);
    const int n = (int)(t / p.P);
    int acc[4] = {0, 0, 0, 0};
    for (int r = 0; r < p.R; ++r) {
      const int ih = op * p.sh - p.pt + r * p.dh;
      if (ih < 0 || ih >= p.H) continue;
      for (int s = 0; s < p.S; ++s) {
        const int iw = oq * p.sw - p.pl_ + s * p.dw;
        if (iw < 0 || iw >= p.W) continue;
        const int8_t* xp = p.x + (((long long)n * p.H + ih) * p.W + iw) * p.ldx + c0;
        if (aligned) {
          const uint32_t xw = ldg_u32(xp);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = c0 + j;
            const uint32_t wv = c < p.C ? (uint32_t)(uint8_t)__ldg(p.w + (long long)c * RS + r * p.S + s) : 0u;
            acc[j] = __dp4a((int)xw, (int)(wv << (8 * j)), acc[j]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = c0 + j;
            if (c < p.C) acc[j] += (int)__ldg(xp + j) * (int)__ldg(p.w + (long long)c * RS + r * p.S + s);
          }
        }
      }
    }
    dw_store4(p, acc, sc, bs, px, c0);
  }
}

}  // namespace pl

using namespace pl;

extern "C" int pl_dwconv2d_i8_nhwc(const int8_t* x, const int8_t* w_crs, void* y,
                                   const pl_conv_desc* d, const pl_epilogue* e, pl_stream_t stream) {
  if (!x || !w_crs || !y || !d || !e) return PL_ERR_ARG;
  if (d->groups != d->C || d->K != d->C) return PL_ERR_ARG;
  if (e->out_dtype != PL_OUT_INT32 && !e->scale) return PL_ERR_ARG;
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->C <= 0 || d->R <= 0 || d->S <= 0 || d->stride_h <= 0 ||
      d->stride_w <= 0 || d->dil_h <= 0 || d->dil_w <= 0 || d->ldx < d->C || d->ldy < d->K)
    return PL_ERR_ARG;
  const int P = pl_conv_out_size(d->H, d->R, d->dil_h, d->pad_t, d->pad_b, d->stride_h);
  const int Q = pl_conv_out_size(d->W, d->S, d->dil_w, d->pad_l, d->pad_r, d->stride_w);
  if (P <= 0 || Q <= 0) return PL_ERR_ARG;
  if ((long long)d->R * d->S * 127 * 127 >= (1 << 22)) return PL_ERR_UNSUPPORTED;  // small_int2float

  DwParams p{};
  p.x = x; p.w = w_crs; p.y = y;
  p.N = d->N; p.H = d->H; p.W = d->W; p.C = d->C; p.ldx = d->ldx; p.ldy = d->ldy; p.P = P; p.Q = Q;
  p.R = d->R; p.S = d->S; p.sh = d->stride_h; p.sw = d->stride_w; p.pt = d->pad_t; p.pl_ = d->pad_l;
  p.dh = d->dil_h; p.dw = d->dil_w;
  p.CG = (d->C + 3) / 4;
  p.npix = (long long)d->N * P * Q;
  p.npairs = (long long)d->N * P * ((Q + 1) / 2);
  p.epi = make_epi(e, d->C);
  if (e->residual) return PL_ERR_UNSUPPORTED;

  int CGB = 1;
  while (CGB < p.CG && CGB < 256) CGB <<= 1;
  const int pper = 256 / CGB;
  const int gy = (p.CG + CGB - 1) / CGB;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) & 3) == 0) && (d->ldx % 4 == 0);
  const bool fast = aligned && d->R == 3 && d->S == 3 && d->dil_h == 1 && d->dil_w == 1 &&
                    d->stride_h == d->stride_w && (d->stride_h == 1 || d->stride_h == 2);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long items = fast ? p.npairs : p.npix;
  long long gx = (items + pper - 1) / pper;
  const long long cap = (long long)sms * 8 / gy > 0 ? (long long)sms * 8 / gy : 1;
  if (gx > cap) gx = cap;
  dim3 grid((unsigned)gx, (unsigned)gy);
  if (fast && d->stride_h == 1)
    dw3x3_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(p, CGB);
  else if (fast)
    dw3x3_kernel<2><<<grid, 256, 0, (cudaStream_t)stream>>>(p, CGB);
  else
    dw_generic_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p, CGB, aligned ? 1 : 0);
  ++g_launches;
  return launch_status();
}
