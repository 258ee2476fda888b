// dwconv_i8.cu -- depthwise int8 conv over NHWC (groups == C == K).
//
// Bandwidth-bound (9 MAC per 2 bytes): no tensor cores.  The reference has no
// kCUDA int8 depthwise at all (cuDNN int8 path ignores groups,
// lite/backends/cuda/math/cudnn_conv.cc:320-585); semantics follow the ARM
// kernels lite/backends/arm/math/conv3x3s{1,2}_depthwise_int8.cc /
// conv5x5s{1,2}_depthwise_int8.cc: exact int32 accumulation, then the common
// epilogue (common.cuh).  Filter layout as the reference stores it: [C][R][S].
//
// Fast path (3x3, stride 1 or 2, dilation 1): see dw3x3s1_kernel / dw3x3s2_kernel.
// Generic path (any R,S,stride,dilation): one dp4a per MAC against a byte-masked
// weight word.
#include <cuda.h>
#include <cstdlib>
#include <cudaTypedefs.h>

#include <type_traits>

#include "common.cuh"

namespace pl {

struct DwParams {
  const int8_t* x;
  const int8_t* w;  // [C][R][S]
  void* y;
  int N, H, W, C, ldx, ldy, P, Q;
  int R, S, sh, sw, pt, pl_, dh, dw;
  int CG;       // channel groups of 4
  long long npix;    // N * P * Q           (generic path)
  long long items;   // fast path: N * bands * strips * CG work items
  int strips, bands;
  FastDiv dCG, dStrips, dBands;
  int ldxw, rowpitch_w;  // pixel / row pitch of the input in 32-bit words
  uint32_t hi_bits;  // float bits of the upper clamp (fast int8 relu epilogue)
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
// fast path: 3x3, dilation 1, stride 1 or 2.
//
// One thread = one work item = 4 channels x (RB output rows) x (T output columns),
// T = 4 (stride 1) or 2 (stride 2).  Consecutive lanes take consecutive channel quads,
// so every warp-level 32-bit load/store covers up to 128 contiguous bytes of a pixel.
// All input words of the item are requested up front (memory-level parallelism comes
// from the loads in flight, not from occupancy); each 4-pixel row segment is byte
// transposed ONCE (8 PRMT) and then reused by the three filter rows, so a 3-tap filter
// row costs one dp4a per output and channel.  The accumulators start at the float
// magic 0x4B400000, which makes int -> float a single (packed) add.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t dw_pk2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void dw_upk2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t dw_ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t dw_fadd2_rn(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t dw_fadd2_rm(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

static constexpr int kDwMagic = 0x4B400000;  // float 12582912 = 1.5 * 2^23

// 4x4 byte transpose: t[c] = {x0.c, x1.c, x2.c, x3.c}
__device__ __forceinline__ void dw_transpose(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t (&t)[4]) {
  const uint32_t a = prmt(x0, x1, 0x5140), b = prmt(x2, x3, 0x5140);
  const uint32_t c = prmt(x0, x1, 0x7362), d = prmt(x2, x3, 0x7362);
  t[0] = prmt(a, b, 0x5410);
  t[1] = prmt(a, b, 0x7632);
  t[2] = prmt(c, d, 0x5410);
  t[3] = prmt(c, d, 0x7632);
}

// epilogue of 4 channels of one output pixel (accumulators carry the magic offset).
// FAST: int8 out, ARM rounding, relu / relu6 (hi_bits = float bits of the upper clamp).
template <bool FAST>
__device__ __forceinline__ void dw_finish4(const DwParams& p, const int (&acc)[4], const float4 sc, const float4 bs,
                                           uint32_t hi_bits, uint32_t pix, int c0) {
  if (FAST) {
    const uint64_t off = dw_pk2(-12582912.f, -12582912.f);
    const uint64_t f01 = dw_fadd2_rn(dw_pk2(__int_as_float(acc[0]), __int_as_float(acc[1])), off);
    const uint64_t f23 = dw_fadd2_rn(dw_pk2(__int_as_float(acc[2]), __int_as_float(acc[3])), off);
    const uint64_t v01 = dw_ffma2(f01, dw_pk2(sc.x, sc.y), dw_pk2(bs.x, bs.y));
    const uint64_t v23 = dw_ffma2(f23, dw_pk2(sc.z, sc.w), dw_pk2(bs.z, bs.w));
    float v0, v1, v2, v3;
    dw_upk2(v01, v0, v1);
    dw_upk2(v23, v2, v3);
    // clamp to [0, hi] on the float bit pattern, then floor(v + 0.5) from byte 1 of
    // RD(v + 32768.5) (see igemm_i8.cu:quad_i8_relu)
    const int c0_ = __vimin_s32_relu(__float_as_int(v0), (int)hi_bits);
    const int c1_ = __vimin_s32_relu(__float_as_int(v1), (int)hi_bits);
    const int c2_ = __vimin_s32_relu(__float_as_int(v2), (int)hi_bits);
    const int c3_ = __vimin_s32_relu(__float_as_int(v3), (int)hi_bits);
    const uint64_t big = dw_pk2(32768.5f, 32768.5f);
    const uint64_t u01 = dw_fadd2_rm(dw_pk2(__int_as_float(c0_), __int_as_float(c1_)), big);
    const uint64_t u23 = dw_fadd2_rm(dw_pk2(__int_as_float(c2_), __int_as_float(c3_)), big);
    const uint32_t p01 = prmt((uint32_t)u01, (uint32_t)(u01 >> 32), 0x0051);
    const uint32_t p23 = prmt((uint32_t)u23, (uint32_t)(u23 >> 32), 0x0051);
    // FAST launches have ldy % 4 == 0: the output is addressed in 32-bit words
    reinterpret_cast<uint32_t*>(reinterpret_cast<int8_t*>(p.y) + c0)[pix * (uint32_t)(p.ldy >> 2)] = prmt(p01, p23, 0x5410);
  } else {
    const int raw[4] = {acc[0] - kDwMagic, acc[1] - kDwMagic, acc[2] - kDwMagic, acc[3] - kDwMagic};
    const float s4[4] = {sc.x, sc.y, sc.z, sc.w}, b4[4] = {bs.x, bs.y, bs.z, bs.w};
    dw_store4(p, raw, s4, b4, (long long)pix, c0);
  }
}

// FAST epilogue of 4 channels of one output pixel straight to a word pointer (the staged kernel keeps a row
// pointer per item instead of re-deriving pixel * pitch + channel for every pixel)
__device__ __forceinline__ void dw_finish4_ptr(const int (&acc)[4], const float4 sc, const float4 bs, uint32_t hi_bits,
                                               uint32_t* dst, bool ok) {
  const uint64_t off = dw_pk2(-12582912.f, -12582912.f);
  const uint64_t f01 = dw_fadd2_rn(dw_pk2(__int_as_float(acc[0]), __int_as_float(acc[1])), off);
  const uint64_t f23 = dw_fadd2_rn(dw_pk2(__int_as_float(acc[2]), __int_as_float(acc[3])), off);
  const uint64_t v01 = dw_ffma2(f01, dw_pk2(sc.x, sc.y), dw_pk2(bs.x, bs.y));
  const uint64_t v23 = dw_ffma2(f23, dw_pk2(sc.z, sc.w), dw_pk2(bs.z, bs.w));
  float v0, v1, v2, v3;
  dw_upk2(v01, v0, v1);
  dw_upk2(v23, v2, v3);
  const int c0_ = __vimin_s32_relu(__float_as_int(v0), (int)hi_bits);
  const int c1_ = __vimin_s32_relu(__float_as_int(v1), (int)hi_bits);
  const int c2_ = __vimin_s32_relu(__float_as_int(v2), (int)hi_bits);
  const int c3_ = __vimin_s32_relu(__float_as_int(v3), (int)hi_bits);
  const uint64_t big = dw_pk2(32768.5f, 32768.5f);
  const uint64_t u01 = dw_fadd2_rm(dw_pk2(__int_as_float(c0_), __int_as_float(c1_)), big);
  const uint64_t u23 = dw_fadd2_rm(dw_pk2(__int_as_float(c2_), __int_as_float(c3_)), big);
  const uint32_t p01 = prmt((uint32_t)u01, (uint32_t)(u01 >> 32), 0x0051);
  const uint32_t p23 = prmt((uint32_t)u23, (uint32_t)(u23 >> 32), 0x0051);
  const uint32_t v = prmt(p01, p23, 0x5410);
  if (ok) *dst = v;  // a predicated store: the math above stays branch-free
}

struct DwItem {
  int n, p0, q0, c0;
};

__device__ __forceinline__ DwItem dw_item(const DwParams& p, uint32_t idx, int RB, int T) {
  DwItem it;
  uint32_t rest = fdiv(idx, p.dCG);
  it.c0 = (int)(idx - rest * p.dCG.d) * 4;
  uint32_t t = fdiv(rest, p.dStrips);
  it.q0 = (int)(rest - t * p.dStrips.d) * T;
  rest = t;
  t = fdiv(rest, p.dBands);
  it.p0 = (int)(rest - t * p.dBands.d) * RB;
  it.n = (int)t;
  return it;
}

// filter words of 4 channels: the reference layout [C][3][3] makes them 36 contiguous,
// 4-byte aligned bytes.  wa[r][j] = {w0, w1, w2, 0} of channel c0+j, filter row r.
__device__ __forceinline__ void dw_load_filter(const DwParams& p, int c0, uint32_t (&wa)[3][4]) {
  if (c0 + 4 <= p.C && (reinterpret_cast<uintptr_t>(p.w) & 3) == 0) {
    uint32_t w[10];
#pragma unroll
    for (int i = 0; i < 9; ++i) w[i] = __ldg(reinterpret_cast<const uint32_t*>(p.w + (long long)c0 * 9) + i);
    w[9] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int byte = j * 9 + r * 3, k = byte >> 2, sh = byte & 3;
        // bytes sh, sh+1, sh+2 of the pair (w[k], w[k+1])
        const uint32_t sel = (uint32_t)(sh | ((sh + 1) << 4) | ((sh + 2) << 8) | ((sh + 3) << 12));
        wa[r][j] = prmt(w[k], w[k + 1], sel) & 0x00ffffffu;
      }
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        uint32_t v = 0;
        if (c0 + j < p.C) {
          const int8_t* wp = p.w + (long long)(c0 + j) * 9 + r * 3;
          v = (uint32_t)(uint8_t)__ldg(wp) | ((uint32_t)(uint8_t)__ldg(wp + 1) << 8) |
              ((uint32_t)(uint8_t)__ldg(wp + 2) << 16);
        }
        wa[r][j] = v;
      }
  }
}

__device__ __forceinline__ void dw_load_scale_bias(const DwParams& p, int c0, float4& sc, float4& bs) {
  float s[4], b[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool ok = c0 + j < p.C;
    s[j] = (ok && p.epi.scale) ? __ldg(p.epi.scale + c0 + j) : 0.f;
    b[j] = (ok && p.epi.bias) ? __ldg(p.epi.bias + c0 + j) : 0.f;
  }
  sc = make_float4(s[0], s[1], s[2], s[3]);
  bs = make_float4(b[0], b[1], b[2], b[3]);
}

// window geometry shared by both strides: clamped word offsets of the NC window columns
// plus an all-ones / all-zeros mask per column (zero padding = load a valid neighbour,
// then AND it away; one LOP3 per word also applies the row mask)
template <int NC>
__device__ __forceinline__ void dw_columns(const DwParams& p, const uint32_t* img, int iw0,
                                           const uint32_t* (&cp)[NC], uint32_t (&cm)[NC]) {
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const int iw = iw0 + k;
    const int c = min(max(iw, 0), p.W - 1);
    cp[k] = img + (uint32_t)(c * p.ldxw);
    cm[k] = (iw == c) ? 0xffffffffu : 0u;
  }
}

// ---- stride 1: T = 4 output columns, RB output rows, window (RB+2) x 6 words ----
template <int RB, bool FAST>
__global__ void __launch_bounds__(256, RB <= 2 ? 4 : 2) dw3x3s1_kernel(const DwParams p) {
  pdl_launch_dependents();
  const uint32_t idx = blockIdx.x * 256u + threadIdx.x;
  if (idx >= (uint32_t)p.items) return;
  const DwItem it = dw_item(p, idx, RB, 4);
  const int c0 = it.c0;
  const int ih0 = it.p0 - p.pt;
  const uint32_t* img = reinterpret_cast<const uint32_t*>(p.x + c0) + (uint32_t)(it.n * p.H) * (uint32_t)p.rowpitch_w;
  const uint32_t* cp[6];
  uint32_t cm[6];
  dw_columns<6>(p, img, it.q0 - p.pl_, cp, cm);
  uint32_t wa[3][4];
  dw_load_filter(p, c0, wa);
  pdl_wait();  // the filter is constant; the input comes from the previous kernel
  int acc[RB][4][4];
#pragma unroll
  for (int o = 0; o < RB; ++o)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[o][q][j] = kDwMagic;
#pragma unroll
  for (int i = 0; i < RB + 2; ++i) {
    const int ih = ih0 + i;
    const int ihc = min(max(ih, 0), p.H - 1);
    const uint32_t rm = (ih == ihc) ? 0xffffffffu : 0u;
    const uint32_t roff = (uint32_t)(ihc * p.rowpitch_w);
    uint32_t xw[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) xw[k] = __ldg(cp[k] + roff) & cm[k] & rm;
    uint32_t ta[4], tb[4];
    dw_transpose(xw[0], xw[1], xw[2], xw[3], ta);  // columns 0..3 -> outputs 0, 1
    dw_transpose(xw[2], xw[3], xw[4], xw[5], tb);  // columns 2..5 -> outputs 2, 3
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int o = i - r;
      if (o < 0 || o >= RB) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int w0 = (int)wa[r][j], w1 = (int)(wa[r][j] << 8);
        acc[o][0][j] = __dp4a((int)ta[j], w0, acc[o][0][j]);
        acc[o][1][j] = __dp4a((int)ta[j], w1, acc[o][1][j]);
        acc[o][2][j] = __dp4a((int)tb[j], w0, acc[o][2][j]);
        acc[o][3][j] = __dp4a((int)tb[j], w1, acc[o][3][j]);
      }
    }
  }
  float4 sc, bs;
  dw_load_scale_bias(p, c0, sc, bs);
#pragma unroll
  for (int o = 0; o < RB; ++o) {
    const int op = it.p0 + o;
    if (op >= p.P) break;
    const uint32_t pix = (uint32_t)((it.n * p.P + op) * p.Q + it.q0);
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (it.q0 + q < p.Q) dw_finish4<FAST>(p, acc[o][q], sc, bs, p.hi_bits, pix + q, c0);
  }
}

// ---- stride 1, vertical streaming: T = 4 output columns, a band of BR output rows per thread ----
// The block-of-rows kernel above re-does the per-item work (item decode, filter unpack, column
// setup, scale / bias loads: ~120 instructions) for every 64 output bytes and is instruction-issue
// bound (ncu: 17 thread-instructions per output byte, 63% issue utilisation at 2.3 TB/s).  Here a
// thread walks down its 4-column strip: every input row is loaded and byte-transposed ONCE, feeds
// the three output rows in flight (accumulator slots rotate through a 3x unrolled loop), the next
// row's loads are issued before the current row's math, and each finished output row leaves through
// the packed epilogue.  ~10 instructions per output byte.
template <bool FAST>
__global__ void __launch_bounds__(256, 2) dw3x3s1v_kernel(const DwParams p, const int BR) {
  pdl_launch_dependents();
  const uint32_t idx = blockIdx.x * 256u + threadIdx.x;
  if (idx >= (uint32_t)p.items) return;
  const DwItem it = dw_item(p, idx, BR, 4);
  const int c0 = it.c0;
  const int ih0 = it.p0 - p.pt;
  const uint32_t* img = reinterpret_cast<const uint32_t*>(p.x + c0) + (uint32_t)(it.n * p.H) * (uint32_t)p.rowpitch_w;
  const uint32_t* cp[6];
  uint32_t cm[6];
  dw_columns<6>(p, img, it.q0 - p.pl_, cp, cm);
  uint32_t wa[3][4], wb[3][4];
  dw_load_filter(p, c0, wa);
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int j = 0; j < 4; ++j) wb[r][j] = wa[r][j] << 8;
  float4 sc, bs;
  dw_load_scale_bias(p, c0, sc, bs);
  pdl_wait();  // filter / scales are constant; the input comes from the previous kernel
  int acc[3][4][4];
#pragma unroll
  for (int o = 0; o < 3; ++o)
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[o][q][j] = kDwMagic;
  const int rows = BR + 2;
  auto load_row = [&](int i, uint32_t (&d)[6]) {
    const int ih = ih0 + i;
    const int ihc = min(max(ih, 0), p.H - 1);
    const uint32_t rm = (ih == ihc) ? 0xffffffffu : 0u;
    const uint32_t roff = (uint32_t)(ihc * p.rowpitch_w);
#pragma unroll
    for (int k = 0; k < 6; ++k) d[k] = __ldg(cp[k] + roff) & cm[k] & rm;
  };
  uint32_t xw[6];
  load_row(0, xw);
  const uint32_t pix0 = (uint32_t)((it.n * p.P + it.p0) * p.Q + it.q0);
  for (int i = 0; i < rows; i += 3) {
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int ii = i + u;
      if (ii >= rows) break;
      uint32_t nx[6];
      if (ii + 1 < rows) load_row(ii + 1, nx);
      uint32_t ta[4], tb[4];
      dw_transpose(xw[0], xw[1], xw[2], xw[3], ta);  // columns 0..3 -> outputs 0, 1
      dw_transpose(xw[2], xw[3], xw[4], xw[5], tb);  // columns 2..5 -> outputs 2, 3
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int o = ii - r;               // output row (relative to the band) this input row feeds with filter row r
        constexpr int kSlotBase = 3;
        const int slot = (u - r + kSlotBase) % 3;  // compile-time: u and r are unrolled
        if (o >= 0 && o < BR) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            acc[slot][0][j] = __dp4a((int)ta[j], (int)wa[r][j], acc[slot][0][j]);
            acc[slot][1][j] = __dp4a((int)ta[j], (int)wb[r][j], acc[slot][1][j]);
            acc[slot][2][j] = __dp4a((int)tb[j], (int)wa[r][j], acc[slot][2][j]);
            acc[slot][3][j] = __dp4a((int)tb[j], (int)wb[r][j], acc[slot][3][j]);
          }
        }
      }
      if (ii >= 2) {  // output row ii - 2 has all three filter rows
        const int o = ii - 2;
        constexpr int dummy = 0;
        (void)dummy;
        const int slot = (u + 1) % 3;
        if (it.p0 + o < p.P) {
          const uint32_t pix = pix0 + (uint32_t)(o * p.Q);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (it.q0 + q < p.Q) dw_finish4<FAST>(p, acc[slot][q], sc, bs, p.hi_bits, pix + q, c0);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[slot][q][j] = kDwMagic;
      }
#pragma unroll
      for (int k = 0; k < 6; ++k) xw[k] = nx[k];
    }
  }
}

// ---- stride 2: T = 2 output columns, RB output rows, window (2*RB+1) x 5 words ----
template <int RB, bool FAST>
__global__ void __launch_bounds__(256, 4) dw3x3s2_kernel(const DwParams p) {
  pdl_launch_dependents();
  const uint32_t idx = blockIdx.x * 256u + threadIdx.x;
  if (idx >= (uint32_t)p.items) return;
  const DwItem it = dw_item(p, idx, RB, 2);
  const int c0 = it.c0;
  const int ih0 = it.p0 * 2 - p.pt;
  const uint32_t* img = reinterpret_cast<const uint32_t*>(p.x + c0) + (uint32_t)(it.n * p.H) * (uint32_t)p.rowpitch_w;
  const uint32_t* cp[5];
  uint32_t cm[5];
  dw_columns<5>(p, img, it.q0 * 2 - p.pl_, cp, cm);
  uint32_t wa[3][4];
  dw_load_filter(p, c0, wa);
  pdl_wait();
  constexpr int NR = 2 * RB + 1;
  int acc[RB][2][4];
#pragma unroll
  for (int o = 0; o < RB; ++o)
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[o][q][j] = kDwMagic;
#pragma unroll
  for (int i = 0; i < NR; ++i) {
    const int ih = ih0 + i;
    const int ihc = min(max(ih, 0), p.H - 1);
    const uint32_t rm = (ih == ihc) ? 0xffffffffu : 0u;
    const uint32_t roff = (uint32_t)(ihc * p.rowpitch_w);
    uint32_t xw[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) xw[k] = __ldg(cp[k] + roff) & cm[k] & rm;
    uint32_t ta[4];
    dw_transpose(xw[0], xw[1], xw[2], xw[3], ta);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      if (((i - r) & 1) != 0 || i - r < 0) continue;
      const int o = (i - r) >> 1;
      if (o >= RB) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // output 0: columns 0,1,2; output 1: columns 2,3 of the transposed word + column 4
        acc[o][0][j] = __dp4a((int)ta[j], (int)wa[r][j], acc[o][0][j]);
        acc[o][1][j] = __dp4a((int)ta[j], (int)(wa[r][j] << 16), acc[o][1][j]);
        acc[o][1][j] = __dp4a((int)xw[4], (int)(((wa[r][j] >> 16) & 0xffu) << (8 * j)), acc[o][1][j]);
      }
    }
  }
  float4 sc, bs;
  dw_load_scale_bias(p, c0, sc, bs);
#pragma unroll
  for (int o = 0; o < RB; ++o) {
    const int op = it.p0 + o;
    if (op >= p.P) break;
    const uint32_t pix = (uint32_t)((it.n * p.P + op) * p.Q + it.q0);
#pragma unroll
    for (int q = 0; q < 2; ++q)
      if (it.q0 + q < p.Q) dw_finish4<FAST>(p, acc[o][q], sc, bs, p.hi_bits, pix + q, c0);
  }
}

// ---------------------------------------------------------------------------
// 3x3 stride 1 / 2, shared-memory staged (the default fast path since round 2)
//
// The register-window kernels above are instruction-issue bound (ncu: 17 thread-instructions per output
// byte, 63 % issue utilisation at 2.3 TB/s): per-item decode, clamped addresses, row / column masks and
// 32-bit global loads.  Here a persistent CTA walks tiles of TR output rows x TQ output columns x CC
// channels.  One 4-D TMA box per tile brings the input rows WITH their halo into shared memory
// (hardware zero fill = the padding: no masks, no clamps), a ring of 2-4 boxes per CTA keeps >= 64 KB
// of loads in flight per SM whatever the thread count, and every thread streams DOWN its strip of 4 (2)
// output columns x 4 channels: each input row is read from shared memory and byte-transposed once,
// feeds the three output rows in flight (accumulator slots rotate through the unrolled loop), and each
// finished output row leaves through the packed epilogue.  ~8 thread-instructions per output byte.
// ---------------------------------------------------------------------------
struct DwTile {
  int TR, strips, quads, CC, BW, BH;
  int ctiles, rtiles, qtiles, tiles;
  int rev_last;     // tiles - 1 when the tile list is walked last-to-first (pl_set_tile_order), else -1
  int l2_hint;      // input boxes are loaded with L2::evict_first (PL_L2_HINTS bit 4)
  int stages;
  int groups, RG;   // row groups of a tile (thread = strip x quad x group) and output rows per group
  int merged;       // the tensor map is {W * C / 8 (8-byte elements), H, N}: whole pixel rows are contiguous
  uint32_t stage_bytes, stage_pitch;
  FastDiv dC, dQ, dR, dQuads, dItems;
};

static constexpr int kDwStagesMax = 4;

template <int STRIDE>
__device__ __forceinline__ void dw_tile_coords(const DwParams& p, const DwTile& t, uint32_t tile, int& n, int& p0,
                                               int& q0, int& cbase) {
  uint32_t rest = fdiv(tile, t.dC);
  cbase = (int)(tile - rest * t.dC.d) * t.CC;
  uint32_t u = fdiv(rest, t.dQ);
  q0 = (int)(rest - u * t.dQ.d) * t.strips * (STRIDE == 1 ? 4 : 2);
  rest = u;
  u = fdiv(rest, t.dR);
  p0 = (int)(rest - u * t.dR.d) * t.TR;
  n = (int)u;
}

#ifndef PL_DW_WARPS
#define PL_DW_WARPS 7
#endif
static constexpr int kDwConsumerWarps = PL_DW_WARPS;
static constexpr int kDwThreads = 32 * (kDwConsumerWarps + 1);  // + the TMA producer warp
// (8 warps = 2 per scheduler: two CTAs of up to 128 registers per thread fit the 16 K registers of an SMSP;
// 9 warps did not -- ncu launch__occupancy_limit_registers 1)

#ifndef PL_DW_CTAS
#define PL_DW_CTAS 2  // resident CTAs per SM the register budget is cut for (experiment: 3 -> 80 registers)
#endif
// CCW > 0: the channel chunk (in words) is a compile-time constant, so the six (five) window loads of a row take
// immediate offsets
template <int STRIDE, bool FAST, int CCW = 0>
__global__ void __launch_bounds__(kDwThreads, PL_DW_CTAS)
dw3x3_tma_kernel(const __grid_constant__ CUtensorMap tx, const DwParams p, const DwTile t) {
  extern __shared__ __align__(128) uint8_t dw_smem[];
  __shared__ __align__(8) uint64_t full_bar[kDwStagesMax], empty_bar[kDwStagesMax];
  pdl_launch_dependents();
  const uint32_t sbase = (smem_u32(dw_smem) + 127u) & ~127u;
  const uint8_t* gbase = dw_smem + (sbase - smem_u32(dw_smem));
  if (threadIdx.x == 0) {
    for (int s = 0; s < t.stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), kDwConsumerWarps);
    }
    fence_barrier_init();
    tma_prefetch_desc(&tx);
  }
  __syncthreads();
  pdl_wait();  // the input comes from the previous kernel
  const int warp = threadIdx.x >> 5;
  if (warp == kDwConsumerWarps) {
    // ---- producer: one lane keeps the ring of input boxes full; no block-wide barrier anywhere ----
    if ((threadIdx.x & 31) == 0) {
      // ring position kept incrementally (a runtime % and / per tile is ~60 instructions)
      int slot = 0;
      uint32_t round = 0;  // how many times the ring has wrapped
      for (uint32_t tile = blockIdx.x; tile < (uint32_t)t.tiles; tile += gridDim.x) {
        if (round > 0) mbar_wait(smem_u32(&empty_bar[slot]), (round - 1u) & 1u);
        int n, p0, q0, cbase;
        dw_tile_coords<STRIDE>(p, t, t.rev_last >= 0 ? (uint32_t)t.rev_last - tile : tile, n, p0, q0, cbase);
        const uint32_t bar = smem_u32(&full_bar[slot]);
        const uint32_t dst = sbase + (uint32_t)slot * t.stage_pitch;
        mbar_arrive_expect_tx(bar, t.stage_bytes);
        // (the input is read once, apart from the halo rows and columns of neighbouring tiles: L2 evict_first, so
        // that the output stays in the cache for the next layer)
        if (t.merged) {  // rows of BW whole pixels: {(W * C) / 8, H, N} in 8-byte elements
          if (t.l2_hint) tma_load_3d_hint(dst, &tx, bar, (q0 * STRIDE - p.pl_) * (p.C >> 3), p0 * STRIDE - p.pt, n, kL2EvictFirst);
          else tma_load_3d(dst, &tx, bar, (q0 * STRIDE - p.pl_) * (p.C >> 3), p0 * STRIDE - p.pt, n);
        } else {
          if (t.l2_hint) tma_load_4d_hint(dst, &tx, bar, cbase, q0 * STRIDE - p.pl_, p0 * STRIDE - p.pt, n, kL2EvictFirst);
          else tma_load_4d(dst, &tx, bar, cbase, q0 * STRIDE - p.pl_, p0 * STRIDE - p.pt, n);
        }
        if (++slot == t.stages) { slot = 0; ++round; }
      }
    }
    return;
  }
  // ---- consumers: a thread owns (row group, strip, channel quad) for the whole kernel.  The grid is a multiple
  // of ctiles and the channel chunk is the fastest tile index, so a CTA always sees the same channel chunk:
  // filter, scale and bias are loaded ONCE (they were re-read per tile: ncu long_scoreboard ~1 per issue) ----
  const int n_items = t.strips * t.quads;
  const uint32_t grp = fdiv(threadIdx.x, t.dItems);
  const int item = (int)threadIdx.x - (int)grp * n_items;
  const uint32_t strip = fdiv((uint32_t)item, t.dQuads);
  const int quad = item - (int)strip * t.quads;
  // (tiles and the grid are multiples of ctiles: the reversed list keeps one channel chunk per CTA as well)
  const uint32_t first = t.rev_last >= 0 ? (uint32_t)t.rev_last - blockIdx.x : blockIdx.x;
  const int cbase0 = (int)(first % (uint32_t)t.ctiles) * t.CC;
  const int c0 = cbase0 + quad * 4;
  const bool active = (int)grp < t.groups && c0 < p.C;
  uint32_t wa[3][4];
  float4 sc, bs;
  if (active) {
    dw_load_filter(p, c0, wa);
    dw_load_scale_bias(p, c0, sc, bs);
  } else {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int j = 0; j < 4; ++j) wa[r][j] = 0;
    sc = bs = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const uint32_t ldyw = (uint32_t)p.ldy >> 2;  // FAST launches: int8 output addressed in 32-bit words
  const uint32_t hi_bits = p.hi_bits;
  const uint32_t colw = CCW > 0 ? (uint32_t)CCW : (uint32_t)t.CC / 4u, roww = (uint32_t)t.BW * colw;
  // word offset of this thread's window inside a stage: first row of its group, first column of its strip
  const uint32_t win_off = (uint32_t)grp * (uint32_t)(t.RG * STRIDE) * roww + strip * 4u * colw + (uint32_t)quad;
  int slot = -1;
  uint32_t parity = 1u;
  for (uint32_t tile = blockIdx.x; tile < (uint32_t)t.tiles; tile += gridDim.x) {
    if (++slot == t.stages) slot = 0;
    if (slot == 0) parity ^= 1u;
    int n, p0, q0, cbase;
    dw_tile_coords<STRIDE>(p, t, t.rev_last >= 0 ? (uint32_t)t.rev_last - tile : tile, n, p0, q0, cbase);
    mbar_wait(smem_u32(&full_bar[slot]), parity);
    const uint8_t* st = gbase + (size_t)slot * t.stage_pitch;
    const int oq0 = q0 + (int)strip * (STRIDE == 1 ? 4 : 2);
    const int pg0 = p0 + (int)grp * t.RG;  // first output row of this thread's group
    int rows_out = min(min(t.RG, t.TR - (int)grp * t.RG), p.P - pg0);
    if (!active || oq0 >= p.Q) rows_out = 0;
    if (rows_out > 0) {
      const uint32_t pix0 = (uint32_t)((n * p.P + pg0) * p.Q + oq0);
      // whole strip inside the image: no per-pixel column checks
      const bool full_strip = oq0 + (STRIDE == 1 ? 4 : 2) <= p.Q;
      // whole strips take a copy of the loop without per-pixel store predicates (most strips are whole: the
      // predicated copy costs ~12 instructions per output row in predicate and descriptor set-up)
      auto run_tile = [&](auto full_tag) {
      constexpr bool FULLS = decltype(full_tag)::value;
      // 32-bit word offsets from the (uniform) base of y: one integer add per pixel instead of a 64-bit pointer
      // chain (the host takes this kernel only when the output has fewer than 2^32 words)
      uint32_t* const ybase = reinterpret_cast<uint32_t*>(p.y);
      uint32_t orow = pix0 * ldyw + ((uint32_t)c0 >> 2);
      const uint32_t orow_step = (uint32_t)p.Q * ldyw;
      auto finish = [&](const int (&acc)[4], int q, uint32_t pix) {
        if (FAST) {
          dw_finish4_ptr(acc, sc, bs, hi_bits, ybase + (orow + (uint32_t)q * ldyw), FULLS || oq0 + q < p.Q);
        } else {
          if (oq0 + q < p.Q) dw_finish4<false>(p, acc, sc, bs, p.hi_bits, pix + q, c0);
        }
      };
      const uint32_t* win = reinterpret_cast<const uint32_t*>(st) + win_off;
      const int magic = kDwMagic;
      if (STRIDE == 1) {
        // input-stationary: three transposed input rows stay in registers (slots rotate through the
        // 3x unrolled loop); every output row is 3 dp4a per output and channel off the float magic, then
        // the packed epilogue -- accumulators live for one pixel only
        uint32_t wb[3][4];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int j = 0; j < 4; ++j) wb[r][j] = wa[r][j] << 8;
        uint32_t T[3][8];
        auto load_row = [&](int ii, uint32_t (&d)[8]) {
          const uint32_t* rp = win + (uint32_t)ii * roww;
          uint32_t xw[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) xw[k] = rp[(uint32_t)k * colw];
          uint32_t ta[4], tb[4];
          dw_transpose(xw[0], xw[1], xw[2], xw[3], ta);  // columns 0..3 -> outputs 0, 1
          dw_transpose(xw[2], xw[3], xw[4], xw[5], tb);  // columns 2..5 -> outputs 2, 3
#pragma unroll
          for (int j = 0; j < 4; ++j) { d[j] = ta[j]; d[4 + j] = tb[j]; }
        };
        load_row(0, T[0]);
        load_row(1, T[1]);
        for (int o = 0; o < rows_out; o += 3) {
#pragma unroll
          for (int u = 0; u < 3; ++u) {
            if (o + u >= rows_out) break;
            load_row(o + u + 2, T[(u + 2) % 3]);
            const uint32_t pix = pix0 + (uint32_t)((o + u) * p.Q);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              int acc[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int tj = (q >> 1) * 4 + j;  // outputs 0, 1 read ta, outputs 2, 3 read tb
                int a_ = __dp4a((int)T[u % 3][tj], (int)((q & 1) ? wb[0][j] : wa[0][j]), magic);
                a_ = __dp4a((int)T[(u + 1) % 3][tj], (int)((q & 1) ? wb[1][j] : wa[1][j]), a_);
                acc[j] = __dp4a((int)T[(u + 2) % 3][tj], (int)((q & 1) ? wb[2][j] : wa[2][j]), a_);
              }
              finish(acc, q, pix);
            }
            orow += orow_step;
          }
        }
      } else {
        // stride 2: window of 5 input columns -> 2 output columns; output m reads input rows 2m, 2m + 1, 2m + 2
        // (row 2m + 2 is carried over as the next output's first row)
        uint32_t wb[3][4], wc[3][4];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            wb[r][j] = wa[r][j] << 16;                                 // taps 0, 1 on transposed columns 2, 3
            wc[r][j] = ((wa[r][j] >> 16) & 0xffu) << (8 * j);          // tap 2 on channel j of column 4
          }
        uint32_t T[3][5];  // [slot][transposed columns 0..3 of channel j | raw column 4]
        auto load_row = [&](int ii, uint32_t (&d)[5]) {
          const uint32_t* rp = win + (uint32_t)ii * roww;
          uint32_t xw[5];
#pragma unroll
          for (int k = 0; k < 5; ++k) xw[k] = rp[(uint32_t)k * colw];
          uint32_t ta[4];
          dw_transpose(xw[0], xw[1], xw[2], xw[3], ta);
#pragma unroll
          for (int j = 0; j < 4; ++j) d[j] = ta[j];
          d[4] = xw[4];
        };
        load_row(0, T[0]);
        for (int m = 0; m < rows_out; m += 2) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            if (m + u >= rows_out) break;
            // slots: u = 0: rows (T0, T1, T2), u = 1: rows (T2, T1, T0)
            const int sa = u == 0 ? 0 : 2, sb_ = 1, sc_ = u == 0 ? 2 : 0;
            load_row(2 * (m + u) + 1, T[sb_]);
            load_row(2 * (m + u) + 2, T[sc_]);
            const uint32_t pix = pix0 + (uint32_t)((m + u) * p.Q);
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              int acc[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                int a_ = magic;
                const int rows3[3] = {sa, sb_, sc_};
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                  if (q == 0) {
                    a_ = __dp4a((int)T[rows3[r]][j], (int)wa[r][j], a_);
                  } else {
                    a_ = __dp4a((int)T[rows3[r]][j], (int)wb[r][j], a_);
                    a_ = __dp4a((int)T[rows3[r]][4], (int)wc[r][j], a_);
                  }
                }
                acc[j] = a_;
              }
              finish(acc, q, pix);
            }
            orow += orow_step;
          }
        }
      }
    
      };
      if (full_strip) run_tile(std::true_type{});
      else run_tile(std::false_type{});
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(smem_u32(&empty_bar[slot]));  // this warp is done with the slot
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

// tensor map of the int8 NHWC input {C, W, H, N} with a box {bc, bw, bh, 1}; no swizzle; out-of-bounds
// box elements (the padding, the channel tail) read as zero
static PFN_cuTensorMapEncodeTiled_v12000 dw_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = [] {
    cudaDriverEntryPointQueryResult q;
    void* f = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(f);
  }();
  return fn;
}
static int dw_encode_input(CUtensorMap* m, const void* base, int C, int W, int H, int N, int ld, int bc, int bw,
                           int bh) {
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)ld, (cuuint64_t)ld * W, (cuuint64_t)ld * W * H};
  cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = dw_encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PL_OK : PL_ERR_DRIVER;
}

// dense thin tensors (ldx == C < 128): {W * C / 8, H, N} in 8-byte elements, box {bw * C / 8, bh, 1}: one box row is a
// whole run of bw pixels (>= 128 bytes) instead of bw rows of C bytes
static int dw_encode_rows(CUtensorMap* m, const void* base, int C, int W, int H, int N, int bw, int bh) {
  cuuint64_t dims[3] = {(cuuint64_t)W * C / 8, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[2] = {(cuuint64_t)W * C, (cuuint64_t)W * C * H};
  cuuint32_t box[3] = {(cuuint32_t)(bw * C / 8), (cuuint32_t)bh, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = dw_encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_INT64, 3, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PL_OK : PL_ERR_DRIVER;
}

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
  p.epi = make_epi(e, d->C);
  if (e->residual) return PL_ERR_UNSUPPORTED;

  const bool aligned = ((reinterpret_cast<uintptr_t>(x) & 3) == 0) && (d->ldx % 4 == 0);
  // the fast kernels address both tensors with 32-bit offsets
  const bool small = (long long)d->N * d->H * d->W * d->ldx < (1ll << 31) &&
                     (long long)d->N * P * Q * d->ldy < (1ll << 31) && (long long)d->N * P * Q * p.CG < (1ll << 31);
  const bool fast = aligned && small && d->R == 3 && d->S == 3 && d->dil_h == 1 && d->dil_w == 1 &&
                    d->stride_h == d->stride_w && (d->stride_h == 1 || d->stride_h == 2);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaStream_t st = (cudaStream_t)stream;
  if (fast) {
    // specialised epilogue: int8 out, ARM rounding, relu / relu6 with a positive clamp,
    // whole channel quads and word-aligned output pixels
    bool fe = e->out_dtype == PL_OUT_INT8 && e->round_mode == PL_ROUND_ARM && e->scale &&
              (e->act == PL_ACT_RELU || e->act == PL_ACT_RELU6) && d->C % 4 == 0 && d->ldy % 4 == 0 &&
              (reinterpret_cast<uintptr_t>(y) & 3) == 0;
    float hi = 127.f;
    if (e->act == PL_ACT_RELU6 && e->act_arg < hi) hi = e->act_arg;
    if (!(hi > 0.f)) fe = false;
    union { float f; uint32_t u; } cv;
    cv.f = hi;
    p.hi_bits = cv.u;
    const int s1 = d->stride_h == 1;
    const int T = s1 ? 4 : 2;
    p.strips = (Q + T - 1) / T;
    // shared-memory staged kernel: needs a TMA-addressable input (16-byte pixels) and 4-byte channel quads
    static const bool no_tma = [] { const char* v = getenv("PL_DW_NO_TMA"); return v && v[0] == '1'; }();
    if (!no_tma && d->ldx % 16 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && dw_encode_fn()) {
      DwTile t{};
      const int strips_all = p.strips;
      const int items_max = 32 * kDwConsumerWarps;
      const int c16 = (d->C + 15) / 16 * 16;
      // TMA costs ~5 cycles per box ROW whatever its length: 32-byte rows (C = 32) run at 6 B/cycle/SM.  So the
      // innermost box extent is >= 128 bytes: a 128-channel chunk, or -- for thinner dense tensors -- whole pixel
      // rows through a map whose innermost dimension is W * C (8-byte elements, at most 256 per box)
      t.merged = (d->C < 128 && d->ldx == d->C && d->C % 16 == 0 && ((long long)d->W * d->C) % 16 == 0) ? 1 : 0;
      t.CC = d->C >= 128 ? 128 : c16;
      t.quads = t.CC / 4;
      t.ctiles = (d->C + t.CC - 1) / t.CC;
      int smax = items_max / t.quads;
      if (smax < 1) smax = 1;
      if (t.merged) {
        const int bw_max = 2048 / d->C;  // 256 elements of 8 bytes
        const int sm_ = s1 ? (bw_max - 2) / 4 : ((bw_max - 1) / 2) / 2;
        if (sm_ < smax) smax = sm_;
      } else {
        const int sm_ = s1 ? (256 - 2) / 4 : ((256 - 1) / 2) / 2;
        if (sm_ < smax) smax = sm_;
      }
      bool ok = smax >= 1;
      if (ok) {
        // balanced column tiles: the fewest tiles that cover the row, equal widths
        t.qtiles = (strips_all + smax - 1) / smax;
        t.strips = (strips_all + t.qtiles - 1) / t.qtiles;
        const int TQ = t.strips * T;
        t.BW = s1 ? TQ + 2 : 2 * TQ + 1;
        const int rowbytes = t.BW * t.CC;
        const int items = t.strips * t.quads;
        t.groups = items_max / items;
        if (t.groups > 4) t.groups = 4;
        if (t.groups < 1) t.groups = 1;
        const int budget = (PL_DW_CTAS == 2 ? 36 : (PL_DW_CTAS == 1 ? 48 : 32)) * 1024;
        int tr = s1 ? budget / rowbytes - 2 : (budget / rowbytes - 1) / 2;
        if (tr > P) tr = P;
        if (tr > 32) tr = 32;
        if (tr < 1) tr = 1;
        auto tiles_for = [&](int r) { return (long long)d->N * ((P + r - 1) / r) * t.qtiles * t.ctiles; };
        while (tr > 4 && tiles_for(tr) < 4ll * sms) --tr;
        if (t.groups > tr) t.groups = tr;
        t.RG = (tr + t.groups - 1) / t.groups;
        t.TR = tr;
        t.BH = s1 ? tr + 2 : 2 * tr + 1;
        t.rtiles = (P + tr - 1) / tr;
        const long long tiles = tiles_for(tr);
        t.stage_bytes = (uint32_t)(rowbytes * t.BH);
        t.stage_pitch = (t.stage_bytes + 127u) / 128u * 128u;
        int stages = (int)(((PL_DW_CTAS == 2 ? 108u : (PL_DW_CTAS == 1 ? 200u : 72u)) * 1024u) / t.stage_pitch);
        if (stages > kDwStagesMax) stages = kDwStagesMax;
        ok = tiles < (1ll << 31) && t.BH <= 256 && stages >= 2 &&
             (long long)d->N * P * Q * (d->ldy / 4) < (1ll << 32);  // 32-bit output word offsets
        if (ok) {
          t.tiles = (int)tiles;
          t.rev_last = next_tile_order_reversed() ? t.tiles - 1 : -1;
          {
            static const int hints = getenv("PL_L2_HINTS") ? atoi(getenv("PL_L2_HINTS")) : 31;
            const long long io_b = (long long)d->N * d->H * d->W * d->C + (long long)d->N * P * Q * d->C;
            static const int min_mb = getenv("PL_L2_HINT_MIN_MB") ? atoi(getenv("PL_L2_HINT_MIN_MB")) : 48;
            t.l2_hint = ((hints & 16) && io_b >= ((long long)min_mb << 20)) ? 1 : 0;
          }
          t.stages = stages;
          t.dC = make_fastdiv((uint32_t)t.ctiles);
          t.dQ = make_fastdiv((uint32_t)t.qtiles);
          t.dR = make_fastdiv((uint32_t)t.rtiles);
          t.dQuads = make_fastdiv((uint32_t)t.quads);
          t.dItems = make_fastdiv((uint32_t)items);
          CUtensorMap tx;
          const int rc = t.merged ? dw_encode_rows(&tx, x, d->C, d->W, d->H, d->N, t.BW, t.BH)
                                  : dw_encode_input(&tx, x, d->C, d->W, d->H, d->N, d->ldx, t.CC, t.BW, t.BH);
          if (rc == PL_OK) {
            const size_t smem = (size_t)t.stages * t.stage_pitch + 128;
            // a multiple of ctiles, so that a CTA keeps its channel chunk (and its filter registers)
            int grid = t.tiles < PL_DW_CTAS * sms ? t.tiles : PL_DW_CTAS * sms;
            grid -= grid % t.ctiles;
            if (grid < t.ctiles) grid = t.ctiles;
            auto go = [&](auto kern) {
              cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)((PL_DW_CTAS == 1 ? 210 : 112) * 1024));
              cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
              launch_pdl(kern, dim3((unsigned)grid), dim3(kDwThreads), smem, st, tx, p, t);
            };
            const int ccw = t.CC / 4;
            if (s1) {
              if (!fe) go(dw3x3_tma_kernel<1, false>);
              else if (ccw == 32) go(dw3x3_tma_kernel<1, true, 32>);
              else if (ccw == 16) go(dw3x3_tma_kernel<1, true, 16>);
              else if (ccw == 8) go(dw3x3_tma_kernel<1, true, 8>);
              else go(dw3x3_tma_kernel<1, true>);
            } else {
              if (!fe) go(dw3x3_tma_kernel<2, false>);
              else if (ccw == 32) go(dw3x3_tma_kernel<2, true, 32>);
              else if (ccw == 16) go(dw3x3_tma_kernel<2, true, 16>);
              else if (ccw == 8) go(dw3x3_tma_kernel<2, true, 8>);
              else go(dw3x3_tma_kernel<2, true>);
            }
            ++g_launches;
            return launch_status();
          }
        }
      }
    }
    // taller bands reuse more rows per thread; keep at least ~2 full waves of threads
    auto items_for = [&](int rb) { return (long long)d->N * ((P + rb - 1) / rb) * p.strips * p.CG; };
    int RB = 2;
    if (s1 && items_for(4) >= (long long)sms * 2048 * 2) RB = 4;
    // stride 1: vertical streaming with the tallest band that still leaves ~2 waves of threads
    // (measured on B200: 1.7 TB/s against 2.5 TB/s for the block-of-rows kernel -- one prefetched row per
    // thread leaves only ~12 KB of loads in flight per SM -- so it stays an experiment: PL_DW_STREAM=1)
    static const bool dw_stream = [] { const char* v = getenv("PL_DW_STREAM"); return v && v[0] == '1'; }();
    bool stream_s1 = s1 && dw_stream;
    if (stream_s1) {
      RB = 4;
      for (int cand : {56, 28, 14, 8}) {
        if (cand <= P && items_for(cand) >= (long long)sms * 512 * 2) { RB = cand; break; }
      }
      if (P < 4) stream_s1 = false, RB = 2;
    }
    p.bands = (P + RB - 1) / RB;
    p.items = items_for(RB);
    p.dCG = make_fastdiv((uint32_t)p.CG);
    p.dStrips = make_fastdiv((uint32_t)p.strips);
    p.dBands = make_fastdiv((uint32_t)p.bands);
    p.ldxw = d->ldx / 4;
    p.rowpitch_w = d->W * p.ldxw;
    const unsigned grid = (unsigned)((p.items + 255) / 256);
    if (stream_s1) {
      if (fe) launch_pdl(dw3x3s1v_kernel<true>, dim3(grid), dim3(256), 0, st, p, RB);
      else launch_pdl(dw3x3s1v_kernel<false>, dim3(grid), dim3(256), 0, st, p, RB);
    } else if (s1) {
      if (RB == 4) {
        if (fe) launch_pdl(dw3x3s1_kernel<4, true>, dim3(grid), dim3(256), 0, st, p);
        else launch_pdl(dw3x3s1_kernel<4, false>, dim3(grid), dim3(256), 0, st, p);
      } else {
        if (fe) launch_pdl(dw3x3s1_kernel<2, true>, dim3(grid), dim3(256), 0, st, p);
        else launch_pdl(dw3x3s1_kernel<2, false>, dim3(grid), dim3(256), 0, st, p);
      }
    } else {
      if (fe) launch_pdl(dw3x3s2_kernel<2, true>, dim3(grid), dim3(256), 0, st, p);
      else launch_pdl(dw3x3s2_kernel<2, false>, dim3(grid), dim3(256), 0, st, p);
    }
  } else {
    int CGB = 1;
    while (CGB < p.CG && CGB < 256) CGB <<= 1;
    const int pper = 256 / CGB;
    const int gy = (p.CG + CGB - 1) / CGB;
    long long gx = (p.npix + pper - 1) / pper;
    const long long cap = (long long)sms * 8 / gy > 0 ? (long long)sms * 8 / gy : 1;
    if (gx > cap) gx = cap;
    dim3 grid((unsigned)gx, (unsigned)gy);
    dw_generic_kernel<<<grid, 256, 0, st>>>(p, CGB, aligned ? 1 : 0);
  }
  ++g_launches;
  return launch_status();
}
