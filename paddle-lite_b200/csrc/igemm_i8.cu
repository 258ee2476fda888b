// igemm_i8.cu -- dense int8 conv (implicit GEMM over NHWC) and fc/matmul GEMM
// on the 5th-gen tensor cores: TMA (tiled or im2col mode) -> swizzled smem ring
// -> tcgen05.mma kind::i8 (M=128, N=BN, K=32/instr) -> int32 accumulators in
// TMEM (double buffered) -> tcgen05.ld -> fused scale/bias/act/requant epilogue.
//
// Replaces lite/backends/cuda/math/cudnn_conv.cc:320-585 (cuDNN int8 conv into
// an fp32 temporary + a second elementwise kernel, activation.cu:91-446) with
// ONE kernel and no temporary.  Semantics of the epilogue: common.cuh.
//
// Warp roles (576 threads, persistent over output tiles):
//   warp 0      TMA producer (one lane)
//   warp 1      TMEM allocator + MMA issuer (one lane)
//   warps 2..17 epilogue; warp w reads TMEM lanes [32*(w%4), +32) and the 16-column
//               units (w-2)/4 + 4i of the tile, 16 columns per tcgen05.ld.  int8
//               outputs are staged in swizzled smem panels and leave by TMA store;
//               the fused residual arrives through a TMA ring fed by warp 0
// The epilogue runs once per output byte and is ALU-issue bound, so the hot
// variants are specialised at compile time (EPI_*): packed FFMA2/FADD2, one
// VIMNMX.RELU for relu+clamp on the float bit pattern, PRMT packing, scale/bias
// staged in shared memory once per tile.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <vector>
#include <cudaTypedefs.h>

#include "common.cuh"

namespace pl {

static constexpr int kBM = 128;            // rows of D per tile (TMEM lanes)
static constexpr int kEpiWarps = 16;                   // TMA-fed launches
static constexpr int kThreads = 64 + 32 * kEpiWarps + 32;  // producer, MMA, 16 epilogue warps, store warp
// thin-input launches: 8 epilogue warps + 8 gather warps (two threads per A row)
static constexpr int kThinEpiWarps = 8;
static constexpr int kThinWarps = 8;
static constexpr int kThinThreads = 64 + 32 * (kThinEpiWarps + kThinWarps) + 32;  // + store warp
static constexpr int kStoreWarp = 18;     // same index in both layouts
static constexpr int kMaxAccStages = 4;   // TMEM accumulators in flight (512 columns / BN, at most 4)
static constexpr int kMaxTmaStages = 12;  // TMA-fed modes
static constexpr int kMaxStages = 28;  // thin mode: 4 KB A stages, the gather runs up to 4 tiles ahead
static constexpr int kMaxResBufs = 6;   // depth of the TMA residual ring
static constexpr int kMaxOutBufs = 8;   // output staging tiles (small tiles: the TMA store's read-out latency spans several tile periods)
static constexpr int kMaxHaloBufs = 4;  // mode 3: input halo tiles in flight (p.halo_bufs of them)

enum { EPI_GENERIC = 0, EPI_I8_RELU = 1, EPI_I8_SIGNED = 2, EPI_I8_RELU_RES = 3 };

struct IgemmParams {
  int M, Nout;
  int m_tiles, n_tiles;
  int BN, BK, stages;
  int num_kb, cblocks;
  int mode;  // 0: A is a [M, K] matrix (tiled TMA); 1: A is NHWC, im2col TMA;
             // 2: thin input (C <= 4, one word per pixel): A tiles are gathered by producer warps
             // 3: halo reuse (stride-1 R x S conv): a tile is `hk` whole output rows of one image on
             //    the padded grid (width WP = Q + S - 1); ONE tiled TMA load per channel block brings
             //    the (hk + R - 1) x WP input pixels, and tap (r, s) is the same smem tile read
             //    through an MMA descriptor that starts (r * WP + s) rows later
  int hk, WP, bands, R, P;   // mode 3
  int halo_m_tiles;          // set before common_setup: tiles along M in halo mode
  uint32_t a_halo_bytes, a_halo_tx;
  uint32_t a_region_bytes;   // bytes of the A stage ring / halo buffers in front of the B ring
  FastDiv dBands, dWP;
  FastDiv dNT;  // n_tiles (set by launch_igemm)
  // tile order hint (pl_set_tile_order): walk the tile list last-to-first, so that a layer starts on the bytes its
  // producer wrote last (still in L2); rev_last = number of tiles - 1, or -1 for the forward order
  int rev_last;
  int l2_hints;  // bit 0: the fused residual is loaded with L2::evict_first, bit 1: the A rows too (one N tile: read once)
  const uint32_t* xw;  // mode 2: input words [N][H][W]
  int thin_r, thin_sw;
  // space-to-depth stem (mode 3 with s2d = 1): the input is [N][HP][WS] 16-byte pixels (2x2 original
  // pixels x 4 channels, zero borders in memory); a tile is one output row (<= 128 pixels); its A
  // operand is s2d_ta raw row spans of (128 + taps - 1) pixels fetched by 1-D bulk copies and read
  // through make_smem_desc_pix16 descriptors
  int s2d, s2d_ta, s2d_kpr;            // tap rows, K = 32 steps per tap row
  uint32_t s2d_span_bytes, s2d_span_pitch;
  long long s2d_row_bytes, s2d_img_bytes;
  const int8_t* xs2d;
  // N-split GEMM fused with its all-gather: the store warp writes every staged tile to the local C
  // and to the same tile of n_peers peer C buffers (NVLink peer memory) through tensor maps that
  // live in global memory -- the transfer overlaps the mainloop tile by tile, no collective call
  // narrow tiles: the 16 epilogue warps split into `epi_groups` groups that take alternate tiles, each
  // warp walking more 16-column units of its own tile (TMEM loads pipelined) -- the fixed per-tile
  // cost of the epilogue chain (waits, fences, arrivals: ~900 cycles) is then paid by 1/groups of the
  // warps per tile instead of all of them
  int epi_groups;
  // long-K GEMMs: a tile is a PAIR of 128-row blocks (256 x BN) that share every B stage -- two MMAs per
  // K step into two TMEM accumulators (all 512 columns at BN = 256).  L2 -> SM bytes per MAC drop by a
  // third (the bound of these launches); the epilogue of a pair no longer overlaps the
  // next mainloop, which long K amortises
  int mpair;
  uint32_t pair_sbase;  // staging offset of a pair's second block: a whole tile (GEMM pairs) or Q rows (s2d row pairs)
  // CTA pairs (kernel template flag CTA2): clusters of two CTAs run MMAs of M = 256 (cta_group::2); every CTA
  // loads its own A rows and HALF of each B block, so the L2 -> SM operand bytes per MAC drop by a third at
  // BN = 256 and the filter of a 3x3 halo layer is fetched once per pair instead of once per CTA.  m_tiles
  // then counts pair rows
  int cta2;
  // tile order of long-K GEMMs (mode 0, streamed B): bands of group_m tile rows are walked column by column,
  // so that one wave of CTAs touches ~sqrt(SMs) A row blocks and B column blocks instead of 2-5 and all of B
  // (8192^3: 24 % of the L2 requests missed to DRAM with the row-major order)
  int group_m;
  // fp32 out, act none / relu, no residual, 16-byte aligned rows: the generic instantiation converts and
  // stores its 16 columns inline (four 16-byte stores per thread) instead of calling store_group16
  int f32_fast;
  // k-block groups (modes 0 / 1, TMA-fed): `kgroup` consecutive ring slots fill under ONE full barrier and are
  // released by ONE commit -- the MMA thread's wait + fence + commit (~450 cycles) is paid once per group instead of
  // once per k-block (it exceeds the MMA time of a 128-byte k-block from BN = 128 down: 4 x 67 cycles)
  int kgroup;
  int store_defer;  // store groups the store warp leaves in flight before it releases a staging buffer (0..4)
  const CUtensorMap* peer_maps;
  int n_peers;
  // N-split GEMM fused with its all-gather through NVSwitch multicast: the store warp copies every staged tile to
  // the MULTICAST address of this rank's column block with multimem.st -- one store leaves the GPU, the switch
  // writes it into every rank's result (own included): 1/(G-1) of the peer-store variant's NVLink egress
  int8_t* mc_out;
  int thin_group;  // thin mode with a resident filter: a ring stage holds all k-blocks of a tile
  int thin_shift;  // window starts `thin_shift` (0 / 1) words left of the first tap: the spare word of the
                   // SW-word window leads, which makes the start even for odd left pads at stride 2
  int thin_pairs;  // every window starts on an even word of an 8-byte aligned row: 8-byte cp.async
  int dbg_flags;  // timing experiments (PL_DBG_FLAGS): 1 skip epilogue body, 2 skip MMA issue, 4 skip B loads, 8 skip A loads
  long long* trace;  // PL_DBG_TRACE: per-role clock64 timeline of CTA 0 (tools/trace_timeline.py)
  int dbg_ashift, dbg_baseoff;  // experiment: MMA A-descriptor starts `dbg_ashift` rows into the stage
  int b_resident;  // modes 2 and 3, one N tile: the whole packed filter stays in smem for the kernel's lifetime
  int halo_bufs;   // mode 3: halo buffers in the A region
  int acc_stages;  // TMEM accumulator ring depth
  int H, W;
  FastDiv dQ, dPQ;
  int Q, PQ, S;
  int sh, sw, pt, pl_, dh, dw;
  uint32_t a_stage_bytes, b_stage_bytes, tx_bytes, tmem_cols, acc_stride;
  void* y;
  int ldy;
  int vec_ok;      // 16-byte vector stores allowed
  int res_vec_ok;  // 16-byte vector loads of the residual allowed
  uint32_t hi_bits;  // float bits of the upper clamp for the relu fast paths
  // int8 fast paths: the tile is staged in swizzled smem panels of PW = min(BN,128)
  // columns and written by TMA (full-line stores, hardware edge clipping); the fused
  // residual arrives the same way, loaded by the producer warp
  int tma_out, tma_res, res_bufs, out_bufs;
  int PW, panels, pw_shift;
  uint32_t swz_mask;                 // 7 / 3 / 1 / 0 for 128 / 64 / 32 / 16-byte panel rows
  uint32_t out_stage_off, res_stage_off;  // byte offsets from the 1024-aligned smem base
  // folded scale / bias in smem: the vectors of ALL n-tiles staged once per CTA (sb_all), or,
  // for very wide outputs, two 256-entry buffers refilled per tile
  uint32_t sb_off;
  int sb_all, sb_n;
  EpiParams epi;
};

// ---- packed fp32x2 helpers (FFMA2 / FADD2 on sm_100) ---------------------------
__device__ __forceinline__ uint64_t pk2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2_rz(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t fadd2_rm(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t fadd2_rn(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint32_t prmt_(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// the same store under a per-lane predicate (no divergence region around the epilogue body for the few padded-grid rows)
__device__ __forceinline__ void st_shared_v4_if(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d, bool ok) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\t@p st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n\t}" ::"r"(addr), "r"(a),
               "r"(b), "r"(c), "r"(d), "r"((uint32_t)ok)
               : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

// 4 accumulators -> 4 packed int8, ARM rounding, activation in {relu, relu6}:
// clamp(v, 0, hi) is ONE integer min-with-relu on the float bit pattern (positive
// floats order like their bits, negative floats are negative ints), then
// round-half-away of a non-negative value = floor(v + 0.5) read from the mantissa of
// RD(v + 32768.5).  Bit-exact with common.cuh:round_sat_i8.
template <bool RES>
__device__ __forceinline__ uint32_t quad_i8_relu(const uint32_t* acc, float4 s, float4 b,
                                                 uint32_t hi_bits, uint32_t res_word, float res_scale) {
  uint64_t v01 = ffma2(pk2(__int2float_rn((int)acc[0]), __int2float_rn((int)acc[1])), pk2(s.x, s.y), pk2(b.x, b.y));
  uint64_t v23 = ffma2(pk2(__int2float_rn((int)acc[2]), __int2float_rn((int)acc[3])), pk2(s.z, s.w), pk2(b.z, b.w));
  if (RES) {
    const uint64_t rs = pk2(res_scale, res_scale);
#ifndef PL_EXP_RES_I2F
    // int8 -> float through the mantissa: 2^23 + (r ^ 0x80) - (2^23 + 128), exact (7 issue slots per quad)
    const uint32_t rx = res_word ^ 0x80808080u;
    const uint32_t magic = 0x4B000000u;
    const uint64_t off = pk2(-8388736.f, -8388736.f);
    const uint64_t r01 = fadd2_rn(pk2(__uint_as_float(prmt_(rx, magic, 0x7650)), __uint_as_float(prmt_(rx, magic, 0x7651))), off);
    const uint64_t r23 = fadd2_rn(pk2(__uint_as_float(prmt_(rx, magic, 0x7652)), __uint_as_float(prmt_(rx, magic, 0x7653))), off);
#else
    // measured alternative: I2F.S8 with a byte selector, 113 instead of 126 issue slots per 16 outputs, but
    // the quarter-rate conversion pipe makes the residual layers 3 % slower (86.2 -> 88.8 us at 56x56)
    const uint64_t r01 = pk2((float)(int8_t)(res_word & 0xffu), (float)(int8_t)((res_word >> 8) & 0xffu));
    const uint64_t r23 = pk2((float)(int8_t)((res_word >> 16) & 0xffu), (float)(int8_t)(res_word >> 24));
#endif
    v01 = ffma2(r01, rs, v01);
    v23 = ffma2(r23, rs, v23);
  }
  float v0, v1, v2, v3;
  upk2(v01, v0, v1);
  upk2(v23, v2, v3);
  const int c0 = __vimin_s32_relu(__float_as_int(v0), (int)hi_bits);
  const int c1 = __vimin_s32_relu(__float_as_int(v1), (int)hi_bits);
  const int c2 = __vimin_s32_relu(__float_as_int(v2), (int)hi_bits);
  const int c3 = __vimin_s32_relu(__float_as_int(v3), (int)hi_bits);
  // round-half-away of v in [0, 127] = floor(v + 0.5): ONE add, rounded down, to
  // 2^15 + 0.5 lands in the binade with ulp 2^-8, so the integer part sits in byte 1 of
  // the result (the exact sum is floored to a multiple of 2^-8, which never crosses an
  // integer)
  const uint64_t big = pk2(32768.5f, 32768.5f);
  const uint64_t u01 = fadd2_rm(pk2(__int_as_float(c0), __int_as_float(c1)), big);
  const uint64_t u23 = fadd2_rm(pk2(__int_as_float(c2), __int_as_float(c3)), big);
  const uint32_t p01 = prmt_((uint32_t)u01, (uint32_t)(u01 >> 32), 0x0051);
  const uint32_t p23 = prmt_((uint32_t)u23, (uint32_t)(u23 >> 32), 0x0051);
  return prmt_(p01, p23, 0x5410);
}

// 4 accumulators -> 4 packed int8, ARM rounding, no activation (signed result)
__device__ __forceinline__ uint32_t quad_i8_signed(const uint32_t* acc, float4 s, float4 b) {
  const uint64_t v01 = ffma2(pk2(__int2float_rn((int)acc[0]), __int2float_rn((int)acc[1])), pk2(s.x, s.y), pk2(b.x, b.y));
  const uint64_t v23 = ffma2(pk2(__int2float_rn((int)acc[2]), __int2float_rn((int)acc[3])), pk2(s.z, s.w), pk2(b.z, b.w));
  float v[4];
  upk2(v01, v[0], v[1]);
  upk2(v23, v[2], v[3]);
  const uint64_t big = pk2(32768.5f, 32768.5f);  // see quad_i8_relu: integer part lands in byte 1
  const uint64_t u01 = fadd2_rm(pk2(fminf(fabsf(v[0]), 127.f), fminf(fabsf(v[1]), 127.f)), big);
  const uint64_t u23 = fadd2_rm(pk2(fminf(fabsf(v[2]), 127.f), fminf(fabsf(v[3]), 127.f)), big);
  // the four magnitudes (byte 1 of each sum) and the four signs (msb of each value, replicated over a byte by the
  // sign mode of PRMT) as packed bytes; bytes are then negated where the sign byte is 0xff: 0x80 - m never borrows
  // (m <= 127), and (0x80 - m) ^ 0x80 = -m mod 256 (16 issue slots per quad instead of 25 with per-lane negation)
  const uint32_t M = prmt_(prmt_((uint32_t)u01, (uint32_t)(u01 >> 32), 0x0051), prmt_((uint32_t)u23, (uint32_t)(u23 >> 32), 0x0051), 0x5410);
  const uint32_t S = prmt_(prmt_(__float_as_uint(v[0]), __float_as_uint(v[1]), 0x00fb),
                           prmt_(__float_as_uint(v[2]), __float_as_uint(v[3]), 0x00fb), 0x5410);
  const uint32_t neg = (0x80808080u - (M & S)) ^ 0x80808080u;  // -m in the negative lanes, 0 elsewhere
  return neg | (M & ~S);
}

// generic element (any dtype / act / rounding / residual)
__device__ __forceinline__ float generic_value(int acc, float s, float b, const EpiParams& e,
                                               const int8_t* res_row, int n, bool n_ok) {
  float t = __fmaf_rn(__int2float_rn(acc), s, b);
  if (e.residual && n_ok) t = __fmaf_rn((float)res_row[n], e.res_scale, t);
  return (e.round_mode == PL_ROUND_CUDA_LEGACY) ? apply_act_legacy(t, e.act, e.act_arg, e.vec4)
                                                : apply_act(t, e.act, e.act_arg);
}

// --------------------------------------------------------------------------
// generic epilogue for one 16-column group held in registers (acc[0..15]): any
// dtype / activation / rounding / residual, ragged edges, unaligned outputs
// --------------------------------------------------------------------------
__device__ __noinline__ void store_group16(const IgemmParams& p, const uint32_t* acc, long long row,
                                           int n_base, const float* s_scale, const float* s_bias) {
  const EpiParams& e = p.epi;
  const bool full = (n_base + 16 <= p.Nout);
  if (e.out_dtype == PL_OUT_INT32) {
    int* out = reinterpret_cast<int*>(p.y) + row * p.ldy + n_base;
    if (full && p.vec_ok) {
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<int4*>(out + j) =
            make_int4((int)acc[j], (int)acc[j + 1], (int)acc[j + 2], (int)acc[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (n_base + j < p.Nout) out[j] = (int)acc[j];
    }
    return;
  }
  const int8_t* res_row = e.residual ? e.residual + row * e.ld_res : nullptr;
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j)
    v[j] = generic_value((int)acc[j], s_scale[j], s_bias[j], e, res_row, n_base + j, n_base + j < p.Nout);
  if (e.out_dtype == PL_OUT_FP32) {
    float* out = reinterpret_cast<float*>(p.y) + row * p.ldy + n_base;
    if (full && p.vec_ok) {
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<float4*>(out + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (n_base + j < p.Nout) out[j] = v[j];
    }
  } else {
    int q[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) q[j] = round_sat_i8(v[j], e.round_mode);
    int8_t* out = reinterpret_cast<int8_t*>(p.y) + row * p.ldy + n_base;
    if (full && p.vec_ok) {
      *reinterpret_cast<uint4*>(out) =
          make_uint4(pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]),
                     pack4_i8(q[8], q[9], q[10], q[11]), pack4_i8(q[12], q[13], q[14], q[15]));
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (n_base + j < p.Nout) out[j] = (int8_t)q[j];
    }
  }
}

// --------------------------------------------------------------------------
// the kernel
// --------------------------------------------------------------------------
// TR / TSW > 0 selects the thin-input mode for an R x (<= TSW words) filter window: four extra
// warps (one thread per A row) gather the window words of every filter row straight from the
// one-word-per-pixel input, all loads of a tile in flight at once, and store them into the
// stage buffers in the 32-byte-swizzled K-major layout the MMA descriptors expect.  One
// k-block (32 bytes) = 32 / (4 * TSW) filter rows.  No im2col matrix ever exists in HBM.
// k-steps of one k-block, unrolled: descriptors advance by 32 bytes (2 x 16-byte units)
template <bool CTA2>
__device__ __forceinline__ void umma_i8_t(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if (CTA2) umma_i8_2sm(d_tmem, da, db, idesc, acc);
  else umma_i8(d_tmem, da, db, idesc, acc);
}
template <bool CTA2>
__device__ __forceinline__ void umma_commit_t(uint32_t bar) {
  if (CTA2) umma_commit_2sm(bar);  // the barrier at this offset in both CTAs of the pair
  else umma_commit(bar);
}
template <int KS, bool CTA2 = false>
__device__ __forceinline__ void issue_kblock(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc,
                                             bool fresh) {
#pragma unroll
  for (int k = 0; k < KS; ++k)
    umma_i8_t<CTA2>(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (uint32_t)(!(fresh && k == 0)));
}
template <bool CTA2 = false>
__device__ __forceinline__ void issue_kblock_n(int ksteps, uint32_t d_tmem, uint64_t da, uint64_t db,
                                               uint32_t idesc, bool fresh) {
  if (ksteps == 4) issue_kblock<4, CTA2>(d_tmem, da, db, idesc, fresh);
  else if (ksteps == 2) issue_kblock<2, CTA2>(d_tmem, da, db, idesc, fresh);
  else issue_kblock<1, CTA2>(d_tmem, da, db, idesc, fresh);
}
// The MMA thread is ONE lane: it issues roughly one instruction every 3 cycles, a taken branch costs ~30, an mbarrier
// wait ~70 even when complete, and the tensor pipe queues only ~8 MMAs behind the running one (tools/sync_probe.cu).
// Whatever the thread does between the last MMA of a tile and the first of the next beyond 8 MMA times drains the
// pipe.  Descriptors of a run are therefore walked with immediate / single-register adds from a base that is opaque
// to the compiler (otherwise the invariant ones get hoisted into vector registers: two R2UR per descriptor and tile).
__device__ __forceinline__ uint32_t opaque_u32(uint32_t v) {
  asm volatile("" : "+r"(v));
  return v;
}
// taps [T0, T1) of a resident 3 x 3 filter: A walks the halo tile (next tap: one pixel = 2 * KS units, next filter
// row: row_pitch), B the resident filter blocks (b_step apart)
template <int KS, bool CTA2, int T0, int T1>
__device__ __forceinline__ void issue_taps3_range(uint32_t d_tmem, uint64_t da0, uint64_t db0, uint32_t idesc,
                                                  uint32_t row_pitch, uint32_t b_step, bool fresh) {
#pragma unroll
  for (int tap = T0; tap < T1; ++tap) {
    const int r = tap / 3, sx = tap - 3 * r;
    issue_kblock<KS, CTA2>(d_tmem, da0 + (uint64_t)(r * row_pitch) + (uint64_t)(sx * 2 * KS), db0 + (uint64_t)(tap * b_step), idesc,
                           fresh && tap == 0);
  }
}

// all MMAs of one space-to-depth tile (TA tap rows x KPR K=32 steps), fully unrolled
template <int TA, int KPR>
__device__ __forceinline__ void issue_s2d(uint32_t d_tmem, uint64_t da0, uint64_t db0, uint32_t idesc,
                                          uint32_t pitch16, uint32_t b_step) {
#pragma unroll
  for (int a = 0; a < TA; ++a)
#pragma unroll
    for (int ks = 0; ks < KPR; ++ks)
      umma_i8(d_tmem, da0 + (uint64_t)(a * pitch16 + 2u * ks), db0 + (uint64_t)((a * KPR + ks) * b_step), idesc,
              (uint32_t)((a | ks) != 0));
}

// ---- the MMA thread's tile loop for a resident 3 x 3 filter (halo mode) ----
// Out of line on purpose: inside the kernel body its unrolled runs changed the code layout of the other single-thread
// role loops (space-to-depth stems +4 %, measured); as a function it has its own registers and sits apart.
struct Taps3Args {
  int tile0, tstride, num_tiles, acc_stages, halo_bufs, cblocks, ksteps;
  uint32_t tmem_base, acc_stride, bar_tempty, bar_tfull, bar_afull, bar_aempty, a_lo0, b_lo0, halo_step, b_step, row_pitch, idesc;
  uint64_t desc_hi;
#ifdef PL_DEBUG_KNOBS
  long long* trace;
#endif
};
#ifdef PL_DEBUG_KNOBS
#define PL_T3TRACE(t, e) do { if (a.trace != nullptr && (t) < 64) a.trace[(64 + (t)) * 8 + (e)] = clock64(); } while (0)
#else
#define PL_T3TRACE(t, e) do { } while (0)
#endif
template <bool CTA2>
__device__ __noinline__ void mma_loop_taps3(const Taps3Args a) {
  // one straight run of 9 * ksteps MMAs per halo tile (and channel block).  The barrier waits of the NEXT unit are
  // taken in the middle of the run, while >= 8 MMAs are queued behind the running one, so the thread reaches the next
  // run with nothing but two commits in between (r2 3x3 64: 52.4 -> 45.1 us at b256)
  const uint32_t b_step9 = 9u * a.b_step;
  int as = 0, ab = 0, lt = 0;
  uint32_t aphase = 0, aph = 0;
  uint32_t a_cur = a.a_lo0;  // descriptor start of the current halo buffer
  if (a.tile0 < a.num_tiles) {
    mbar_wait(a.bar_tempty + 8 * as, aphase ^ 1);
    mbar_wait(a.bar_afull + 8 * ab, aph);
  }
  for (int tile = a.tile0; tile < a.num_tiles; tile += a.tstride, ++lt) {
    PL_T3TRACE(lt, 0);
    const uint32_t d_tmem = a.tmem_base + (uint32_t)as * a.acc_stride;
    int as_n = as + 1;
    uint32_t aphase_n = aphase;
    if (as_n == a.acc_stages) { as_n = 0; aphase_n ^= 1; }
    uint32_t b_cur = opaque_u32(a.b_lo0);
    for (int cb = 0; cb < a.cblocks; ++cb) {
      tc_fence_after();
      if (cb == 0) PL_T3TRACE(lt, 2);
      const uint64_t da0 = a.desc_hi | (uint64_t)a_cur, db0 = a.desc_hi | (uint64_t)b_cur;
      const bool last_cb = cb == a.cblocks - 1;
      int ab_n = ab + 1;
      uint32_t aph_n = aph, a_nxt = a_cur + a.halo_step;
      if (ab_n == a.halo_bufs) { ab_n = 0; aph_n ^= 1; a_nxt = a.a_lo0; }
      const bool more = !last_cb || tile + a.tstride < a.num_tiles;
#define PL_TAPS3(KS_)                                                                                          \
  issue_taps3_range<KS_, CTA2, 0, 5>(d_tmem, da0, db0, a.idesc, a.row_pitch, a.b_step, cb == 0);               \
  if (more && !CTA2) {                                                                                         \
    if (last_cb) mbar_wait(a.bar_tempty + 8 * as_n, aphase_n ^ 1);                                             \
    mbar_wait(a.bar_afull + 8 * ab_n, aph_n);                                                                  \
  }                                                                                                            \
  issue_taps3_range<KS_, CTA2, 5, 9>(d_tmem, da0, db0, a.idesc, a.row_pitch, a.b_step, false);
      if (a.ksteps == 4) { PL_TAPS3(4) }
      else if (a.ksteps == 2) { PL_TAPS3(2) }
      else { PL_TAPS3(1) }
#undef PL_TAPS3
      umma_commit_t<CTA2>(a.bar_aempty + 8 * ab);
      if (last_cb) umma_commit_t<CTA2>(a.bar_tfull + 8 * as);
      if (more && CTA2) {  // pairs are bound by the L2 stream: a wait in mid-run would stall MMAs whose operands are there
        if (last_cb) mbar_wait(a.bar_tempty + 8 * as_n, aphase_n ^ 1);
        mbar_wait(a.bar_afull + 8 * ab_n, aph_n);
      }
      ab = ab_n; aph = aph_n; a_cur = a_nxt;
      b_cur += b_step9;
    }
    PL_T3TRACE(lt, 4);
    as = as_n; aphase = aphase_n;
  }
}

// ---- the MMA thread's tile loop of the space-to-depth stems (out of line like mma_loop_taps3) ----
struct S2dArgs {
  int tile0, tstride, num_tiles, acc_stages, halo_bufs, ta, kpr, skip_mma, rows, bn;
  uint32_t tmem_base, acc_stride, bar_tempty, bar_tfull, bar_afull, bar_aempty, a_lo0, b_lo0, halo_step, b_step, pitch16, idesc;
  uint64_t desc_hi;
#ifdef PL_DEBUG_KNOBS
  long long* trace;
#endif
};
__device__ __noinline__ void mma_loop_s2d(const S2dArgs a) {
  // a tile is TA x KPR <= 8 MMAs: all of them fit the tensor pipe's queue, so the waits of the next tile follow the
  // commits directly and run while the queue drains
  const uint64_t desc_a = make_smem_desc_pix16();
  int as = 0, ab = 0, lt = 0;
  uint32_t aphase = 0, aph = 0;
  uint32_t a_cur = a.a_lo0;
  if (a.tile0 < a.num_tiles) {
    mbar_wait(a.bar_tempty + 8 * as, aphase ^ 1);
    mbar_wait(a.bar_afull + 8 * ab, aph);
  }
  for (int tile = a.tile0; tile < a.num_tiles; tile += a.tstride, ++lt) {
    PL_T3TRACE(lt, 0);
    const uint32_t d_tmem = a.tmem_base + (uint32_t)as * a.acc_stride;
    int as_n = as + 1, ab_n = ab + 1;
    uint32_t aphase_n = aphase, aph_n = aph, a_nxt = a_cur + a.halo_step;
    if (as_n == a.acc_stages) { as_n = 0; aphase_n ^= 1; }
    if (ab_n == a.halo_bufs) { ab_n = 0; aph_n ^= 1; a_nxt = a.a_lo0; }
    tc_fence_after();
    PL_T3TRACE(lt, 2);
    const uint32_t b_cur = opaque_u32(a.b_lo0);
    const uint64_t da0 = desc_a | (uint64_t)a_cur, db0 = a.desc_hi | (uint64_t)b_cur;
    if (a.skip_mma) {
    } else if (a.ta == 4 && a.kpr == 2) {  // 7x7 / p3 stem
      issue_s2d<4, 2>(d_tmem, da0, db0, a.idesc, a.pitch16, a.b_step);
      // row pairs: the second output row one span further down, into the accumulator next door
      if (a.rows == 2) issue_s2d<4, 2>(d_tmem + (uint32_t)a.bn, da0 + (uint64_t)a.pitch16, db0, a.idesc, a.pitch16, a.b_step);
    } else if (a.ta == 2 && a.kpr == 1) {  // 3x3 / p1 stem
      issue_s2d<2, 1>(d_tmem, da0, db0, a.idesc, a.pitch16, a.b_step);
      if (a.rows == 2) issue_s2d<2, 1>(d_tmem + (uint32_t)a.bn, da0 + (uint64_t)a.pitch16, db0, a.idesc, a.pitch16, a.b_step);
    } else {
      uint32_t t_lo = a_cur, bb = b_cur;
      for (int r = 0; r < a.ta; ++r, t_lo += a.pitch16)
        for (int ks = 0; ks < a.kpr; ++ks, bb += a.b_step)  // K = 32: pixels (2 ks, 2 ks + 1) of every row
          umma_i8(d_tmem, desc_a | (uint64_t)(t_lo + 2u * (uint32_t)ks), a.desc_hi | (uint64_t)bb, a.idesc, (uint32_t)((r | ks) != 0));
    }
    umma_commit(a.bar_aempty + 8 * ab);
    umma_commit(a.bar_tfull + 8 * as);
    PL_T3TRACE(lt, 4);
    if (tile + a.tstride < a.num_tiles) {
      mbar_wait(a.bar_tempty + 8 * as_n, aphase_n ^ 1);
      mbar_wait(a.bar_afull + 8 * ab_n, aph_n);
    }
    as = as_n; aphase = aphase_n; ab = ab_n; aph = aph_n; a_cur = a_nxt;
  }
}

// timeline tracing: role r, local tile t (first 64 tiles of CTA 0), event e (0..7)
// Debug knobs (timeline trace, PL_DBG_FLAGS ablations, descriptor experiments) exist only in
// builds made with -DPL_DEBUG_KNOBS (tools/build_variant.sh dbg "-DPL_DEBUG_KNOBS"): even
// never-taken checks in the role loops cost 5-10% on the short-K layers.
#ifdef PL_DEBUG_KNOBS
#define PL_DBG_FLAGS_V p.dbg_flags
#define PL_DBG_ASHIFT_V p.dbg_ashift
#define PL_DBG_BASEOFF_V p.dbg_baseoff
#else
#define PL_DBG_FLAGS_V 0
#define PL_DBG_ASHIFT_V 0
#define PL_DBG_BASEOFF_V 0
#endif
#ifndef PL_DEBUG_KNOBS
#define PL_TRACE(r, t, e) do { } while (0)
#else
#define PL_TRACE(r, t, e)                                                                   \
  do {                                                                                      \
    if (p.trace != nullptr && blockIdx.x == 0 && (t) < 64) p.trace[((r) * 64 + (t)) * 8 + (e)] = clock64(); \
  } while (0)
#endif

// GROUPED is a kernel template flag: the banded order lives in its own instantiations (the large GEMMs);
// sharing one kernel behind a runtime branch cost the conv nets 1.7 %
// CTA2: `tile` indexes a PAIR of 128-row blocks (p.m_tiles counts pair rows); the CTA of cluster rank `rank`
// owns block 2 * tm + rank of it
template <bool GROUPED, bool CTA2 = false>
__device__ __forceinline__ void tile_mn(const IgemmParams& p, int tile, int& tm, int& tn, int rank = 0) {
  if (p.rev_last >= 0) tile = p.rev_last - tile;
  if (GROUPED) {
    const int per_band = p.group_m * p.n_tiles;
    const int band = tile / per_band, within = tile - band * per_band;
    const int left = p.m_tiles - band * p.group_m;
    const int rows = left < p.group_m ? left : p.group_m;
    tn = within / rows;
    tm = band * p.group_m + within - tn * rows;
  } else {
    // (a runtime / and % cost ~40 instructions per tile in each of the 18 warps that walk the tile list)
    tm = (int)fdiv((uint32_t)tile, p.dNT);
    tn = tile - tm * p.n_tiles;
  }
  if (CTA2) tm = 2 * tm + rank;
}

template <int EPI, int TR = 0, int TSW = 0, bool GROUPED = false, bool CTA2 = false>
__global__ void __launch_bounds__(TR > 0 ? kThinThreads : kThreads, 1)
igemm_i8_kernel(const __grid_constant__ CUtensorMap tmap_a,
                const __grid_constant__ CUtensorMap tmap_b,
                const __grid_constant__ CUtensorMap tmap_y,
                const __grid_constant__ CUtensorMap tmap_r, const IgemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int kBarTFull = 2 * kMaxStages, kBarTEmpty = kBarTFull + kMaxAccStages;
  constexpr int kBarRFull = kBarTEmpty + kMaxAccStages, kBarREmpty = kBarRFull + kMaxResBufs;
  constexpr int kBarAFull = kBarREmpty + kMaxResBufs, kBarAEmpty = kBarAFull + kMaxHaloBufs;
  // (`staged` serves the residual ring or the output staging buffers, whichever the launch uses)
  constexpr int kMaxStaged = kMaxResBufs > kMaxOutBufs ? kMaxResBufs : kMaxOutBufs;
  constexpr int kBarBRes = kBarAEmpty + kMaxHaloBufs, kBarStaged = kBarBRes + 1, kBarOFree = kBarStaged + kMaxStaged;
  __shared__ __align__(8) uint64_t bars[kBarOFree + kMaxOutBufs];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + p.a_region_bytes;
  const uint32_t bar_full = smem_u32(&bars[0]);
  const uint32_t bar_empty = smem_u32(&bars[kMaxStages]);
  const uint32_t bar_tfull = smem_u32(&bars[kBarTFull]);
  const uint32_t bar_tempty = smem_u32(&bars[kBarTEmpty]);
  const uint32_t bar_rfull = smem_u32(&bars[kBarRFull]);
  const uint32_t bar_rempty = smem_u32(&bars[kBarREmpty]);
  const uint32_t bar_afull = smem_u32(&bars[kBarAFull]);
  const uint32_t bar_aempty = smem_u32(&bars[kBarAEmpty]);
  const uint32_t bar_bres = smem_u32(&bars[kBarBRes]);
  const uint32_t bar_staged = smem_u32(&bars[kBarStaged]);  // staging buffer written by every epilogue warp
  const uint32_t bar_ofree = smem_u32(&bars[kBarOFree]);    // staging buffer read out by its TMA store
  const uint32_t smem_out = smem_base + p.out_stage_off;
  const uint32_t smem_res = smem_base + p.res_stage_off;
  const uint32_t tile_bytes = (uint32_t)(kBM * p.BN);
  const uint32_t obuf_bytes = tile_bytes << p.mpair;  // output staging buffer: one tile, or the two of a pair
  float* const s_scale = reinterpret_cast<float*>(smem_raw + (smem_base - smem_u32(smem_raw)) + p.sb_off);
  float* const s_bias = s_scale + p.sb_n;

  constexpr int EW = TR > 0 ? kThinEpiWarps : kEpiWarps;  // epilogue warps of this instantiation
  // shfl from lane 0: the compiler then knows the role index is warp-uniform (uniform branches and
  // uniform registers for the role loops)
  // Role index = REVERSED hardware warp index: the SM's issue arbiter favours the highest warp id
  // of an SMSP, so the two single-lane roles every tile waits on (TMA producer = role 0, MMA issuer
  // = role 1) take the top warp ids; with the natural order they queued behind 16 busy epilogue
  // warps (measured: ~300 cycles from `full` to the commit of a 2-MMA k-block).
  const int hw_warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
#ifdef PL_EXP_NATURAL_WARP_ORDER
  const int warp = hw_warp;
#else
  const int warp = kStoreWarp - hw_warp;  // both layouts have kStoreWarp + 1 warps
#endif
  const int lane = threadIdx.x & 31;
  // (debug builds: 4096 = no tiles at all -- prologue + teardown only; 8192 = return at entry; 1024 = no scale / bias
  // staging; 512 = no TMEM allocation, only together with 4096)
  const int num_tiles = (PL_DBG_FLAGS_V & 4096) ? 0 : p.m_tiles * p.n_tiles;
  if (PL_DBG_FLAGS_V & 8192) return;
  // CTA pairs: both CTAs of a cluster walk the same sequence of pair tiles
  const int cta_rank = CTA2 ? (int)cluster_ctarank() : 0;
  const int tile0 = CTA2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tstride = CTA2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
#if defined(PL_EXP_POLL) && defined(PL_DEBUG_KNOBS)
  const bool poll_waits = (PL_DBG_FLAGS_V & 256) != 0;
  auto WAIT = [&](uint32_t bar, uint32_t parity) {
    if (poll_waits) mbar_wait_poll(bar, parity);
    else mbar_wait(bar, parity);
  };
#else
#define WAIT(bar, parity) mbar_wait(bar, parity)
#endif

  pdl_launch_dependents();  // the next layer's CTAs may become resident as ours retire (common.cuh)
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (EPI != EPI_GENERIC) tma_prefetch_desc(&tmap_y);
    if (EPI == EPI_I8_RELU_RES) tma_prefetch_desc(&tmap_r);
    for (int s = 0; s < p.stages; ++s) {
      // TMA arrival (B, and A in the TMA modes) + one arrival per gathered A row
      // TMA arrival (B when streamed, A in the TMA modes) + one arrival per gather thread
      mbar_init(bar_full + 8 * s, TR > 0 ? (p.b_resident ? 0 : 1) + (p.thin_group ? 2 * kBM : kBM) : 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < kMaxAccStages; ++s) {
      mbar_init(bar_tfull + 8 * s, 1);
      // one arrival per epilogue warp of the tile's group (pairs: of both CTAs, on the leader's barrier)
      mbar_init(bar_tempty + 8 * s, (CTA2 ? 2 : 1) * (EW / p.epi_groups));
    }
    for (int s = 0; s < kMaxStaged; ++s) mbar_init(bar_staged + 8 * s, EW / p.epi_groups);
    for (int s = 0; s < kMaxOutBufs; ++s) mbar_init(bar_ofree + 8 * s, 1);
    for (int s = 0; s < kMaxHaloBufs; ++s) {
      mbar_init(bar_afull + 8 * s, 1);
      mbar_init(bar_aempty + 8 * s, 1);
    }
    mbar_init(bar_bres, 1);
    for (int s = 0; s < kMaxResBufs; ++s) {
      mbar_init(bar_rfull + 8 * s, 1);
      // fused residual: the output tile is written IN PLACE over the residual tile, so the buffer
      // is released by the store warp (one arrival) once its TMA store has read it out
      mbar_init(bar_rempty + 8 * s, EPI == EPI_I8_RELU_RES ? 1 : EW);
    }
    fence_barrier_init();
  }
  if (warp == 1 && !(PL_DBG_FLAGS_V & 512)) {
    if (CTA2) {
      tmem_alloc_2sm(smem_u32(&tmem_base_slot), p.tmem_cols);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(smem_u32(&tmem_base_slot), p.tmem_cols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (CTA2) cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  // pair mode: the barriers the LEADER's MMA thread waits on (full, afull, tempty) live in the leader CTA;
  // loads of both CTAs complete on them, the peer's epilogue warps arrive on them remotely
  const uint32_t lead_delta = CTA2 ? mapa_shared(bar_full, 0) - bar_full : 0u;

  // Warps 0 and 1 hand their whole loop nest to ONE lane chosen by elect.sync: the compiler
  // then knows a single thread is active and keeps the loop state and every TMA / MMA operand
  // on the uniform datapath.  Wrapping the loops in `if (lane == 0)` instead makes it
  // re-uniformise each operand with an ELECT + BRA.U.ANY loop: ~85 instructions and ~640
  // cycles per k-block, which bounded every 3x3 layer.
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {  // ONE lane runs the whole loop nest: the compiler keeps it on the uniform datapath
    // all loop state is incremental (no multiplies, no parameter reloads per k-block):
    // this warp's issue latency per k-block bounds the short-K 3x3 layers
    const uint32_t tx = p.tx_bytes, a_bytes = p.a_stage_bytes, b_bytes = p.b_stage_bytes;
    const int stages = p.stages, num_kb = p.num_kb, BK = p.BK, S = p.S, cspan = p.cblocks * p.BK;
    const int mode = p.mode, n_tiles = p.n_tiles, BN = p.BN;
    const int KG = (TR > 0 || p.kgroup < 1) ? 1 : p.kgroup;  // ring slots per barrier group (modes 0 / 1)
    const uint16_t dwv = (uint16_t)p.dw, dhv = (uint16_t)p.dh;
    int stage = 0;
    uint32_t phase = 0;
    uint32_t full_b = bar_full, empty_b = bar_empty, dst_a = smem_a, dst_b = smem_b;
    int rb = 0;
    uint32_t rphase = 0;
    // L2 policies of the residual and the activation loads, chosen once: the loads always take the hinted form (a
    // branch per load in this thread cost as much as the hints gain)
    const uint64_t pol_r = (p.l2_hints & 1) ? kL2EvictFirst : kL2EvictNormal;
    const uint64_t pol_a = (p.l2_hints & 14) ? kL2EvictFirst : kL2EvictNormal;
    if (TR > 0 && p.b_resident) {
      // thin mode, single N tile: every k-block of the filter is loaded exactly once
      {
        mbar_arrive_expect_tx(bar_rfull, (uint32_t)num_kb * (uint32_t)(BN * BK));
        for (int kb = 0; kb < num_kb; ++kb)
          tma_load_2d(smem_b + kb * b_bytes, &tmap_b, bar_rfull, kb * BK, 0);
      }
    }
    // the filter is constant: resident-filter loads above (and in halo mode below) are issued
    // before the wait; activations / residual come from the previous kernels
    const bool bres01 = TR == 0 && mode != 3 && p.b_resident;
    if (bres01) {
      // the grid is a multiple of n_tiles: this CTA only ever sees N tile blockIdx.x % n_tiles, whose
      // whole K extent stays in smem (no per-tile filter re-read from L2, A-only ring stages)
      mbar_arrive_expect_tx(bar_bres, (uint32_t)num_kb * (uint32_t)(BN * BK));
      int tm0_ = 0, tn0_ = 0;
      if (tile0 < num_tiles) tile_mn<GROUPED, CTA2>(p, tile0, tm0_, tn0_, cta_rank);  // (honours the tile order hint)
      const int n0r = tn0_ * BN;
      for (int kb = 0; kb < num_kb; ++kb) tma_load_2d(smem_b + kb * b_bytes, &tmap_b, bar_bres, kb * BK, n0r);
    }
    if (!(TR == 0 && mode == 3)) pdl_wait();
    if (TR == 0 && mode == 3) {
      // halo reuse: one tiled 4-D load per (tile, channel block), then the filter taps
      const int RS = p.R * S, halo_bufs = p.halo_bufs;
      const bool hres = p.b_resident != 0;
      if (hres) {
        // single N tile and the packed filter fits: every filter block is loaded exactly once
        // (channel block major, tap minor: the order the MMA loop walks them)
        // (pairs: every CTA keeps ITS half of the filter rows; both halves complete on the leader's barrier)
        if (!CTA2 || cta_rank == 0) mbar_arrive_expect_tx(bar_bres, (CTA2 ? 2u : 1u) * (uint32_t)num_kb * tx);
        uint32_t d = smem_b;
        for (int cbk = 0; cbk < cspan; cbk += BK)
          for (int tap = 0, kc = cbk; tap < RS; ++tap, kc += cspan, d += b_bytes) {
            if (CTA2) tma_load_2d_2sm(d, &tmap_b, bar_bres + lead_delta, kc, cta_rank * (BN >> 1));
            else tma_load_2d(d, &tmap_b, bar_bres, kc, 0);
          }
      }
      pdl_wait();
      int ab = 0;
      uint32_t aph = 0;
      int lt = 0;
      if (p.s2d) {
        // one output row per tile: s2d_ta contiguous spans, one bulk copy each
        const int ta = p.s2d_ta;
        const uint32_t span = p.s2d_span_bytes, pitch = p.s2d_span_pitch;
        for (int tile = tile0; tile < num_tiles; tile += tstride, ++lt) {
          PL_TRACE(0, lt, 0);
          const int te = p.rev_last >= 0 ? p.rev_last - tile : tile;
          const int img = (int)fdiv((uint32_t)te, p.dBands);
          const int p0 = (te - img * p.bands) * p.hk;  // first output row of the tile (row pairs: hk = 2)
          WAIT(bar_aempty + 8 * ab, aph ^ 1);
          PL_TRACE(0, lt, 1);
          mbar_arrive_expect_tx(bar_afull + 8 * ab, p.a_halo_tx);
          const int8_t* src = p.xs2d + (long long)img * p.s2d_img_bytes + (long long)p0 * p.s2d_row_bytes;
          uint32_t dst = smem_a + ab * p.a_halo_bytes;
          const int spans = ta + p.hk - 1;  // output row r of the tile reads spans r .. r + ta - 1
          for (int a = 0; a < spans; ++a, src += p.s2d_row_bytes, dst += pitch) bulk_load_1d(dst, src, span, bar_afull + 8 * ab);
          PL_TRACE(0, lt, 2);
          if (++ab == halo_bufs) { ab = 0; aph ^= 1; }
        }
      }
      // (HINT: the loads always take the hinted TMA form, with the policy chosen once per launch.  A branch per load
      // cost this thread as much as the hints gain, and a second, plain copy of the loop for small launches moved
      // the code enough to lose the gain as well: profiles/r02_l2.md)
      auto halo_tiles = [&](auto hint_tag) {
      constexpr bool HINT = decltype(hint_tag)::value;
      for (int tile = tile0; tile < num_tiles && !p.s2d; tile += tstride, ++lt) {
        PL_TRACE(0, lt, 0);
        const int te = p.rev_last >= 0 ? p.rev_last - tile : tile;
        const int tq = (int)fdiv((uint32_t)te, p.dNT);
        const int tn = te - tq * n_tiles, tm = CTA2 ? 2 * tq + cta_rank : tq;
        const int n0 = tn * BN;
        // (pairs: the block past the last one reads image N -- all zero fill, nothing stored)
        const int img = (int)fdiv((uint32_t)tm, p.dBands);
        const int p0 = (tm - img * p.bands) * p.hk;
        if (EPI == EPI_I8_RELU_RES) {
          WAIT(bar_rempty + 8 * rb, rphase ^ 1);
          {
            mbar_arrive_expect_tx(bar_rfull + 8 * rb, (uint32_t)(p.hk * p.Q * BN));
            for (int j = 0; j < p.panels; ++j) {
              if (HINT)
                tma_load_4d_hint(smem_res + rb * tile_bytes + j * (kBM * p.PW), &tmap_r, bar_rfull + 8 * rb,
                                 n0 + j * p.PW, 0, p0, img, pol_r);
              else
                tma_load_4d(smem_res + rb * tile_bytes + j * (kBM * p.PW), &tmap_r, bar_rfull + 8 * rb,
                            n0 + j * p.PW, 0, p0, img);
            }
          }
          if (++rb == p.res_bufs) { rb = 0; rphase ^= 1; }
        }
        for (int cbk = 0; cbk < cspan; cbk += BK) {
          WAIT(bar_aempty + 8 * ab, aph ^ 1);
          PL_TRACE(0, lt, 1);
          {
            if (CTA2) {  // both halo tiles complete on the leader's barrier
              if (cta_rank == 0) mbar_arrive_expect_tx(bar_afull + 8 * ab, 2u * p.a_halo_tx);
              tma_load_4d_2sm(smem_a + ab * p.a_halo_bytes, &tmap_a, bar_afull + 8 * ab + lead_delta, cbk, -p.pl_,
                              p0 - p.pt, img);
            } else if (PL_DBG_FLAGS_V & 8) {
              mbar_arrive(bar_afull + 8 * ab);
            } else {
              mbar_arrive_expect_tx(bar_afull + 8 * ab, p.a_halo_tx);
              if (HINT) tma_load_4d_hint(smem_a + ab * p.a_halo_bytes, &tmap_a, bar_afull + 8 * ab, cbk, -p.pl_, p0 - p.pt, img, pol_a);
              else tma_load_4d(smem_a + ab * p.a_halo_bytes, &tmap_a, bar_afull + 8 * ab, cbk, -p.pl_, p0 - p.pt, img);
            }
          }
          PL_TRACE(0, lt, 2);
          if (++ab == halo_bufs) { ab = 0; aph ^= 1; }
          int kc = cbk;  // K coordinate of the filter block: tap * cspan + channel offset
          for (int tap = 0; tap < RS && !hres; ++tap) {
            WAIT(empty_b, phase ^ 1);
            {
              if (CTA2) {  // this CTA's half of the filter block, completing on the leader's barrier
                if (cta_rank == 0) mbar_arrive_expect_tx(full_b, 2u * tx);
                tma_load_2d_2sm(dst_b, &tmap_b, full_b + lead_delta, kc, n0 + cta_rank * (BN >> 1));
              } else if (PL_DBG_FLAGS_V & 4) {
                mbar_arrive(full_b);
              } else {
                mbar_arrive_expect_tx(full_b, tx);
                tma_load_2d(dst_b, &tmap_b, full_b, kc, n0);
              }
            }
            kc += cspan;
            ++stage; full_b += 8; empty_b += 8; dst_b += b_bytes;
            if (stage == stages) { stage = 0; phase ^= 1; full_b = bar_full; empty_b = bar_empty; dst_b = smem_b; }
          }
        }
      }
      };  // halo_tiles
      halo_tiles(std::true_type{});
      if (CTA2) {  // (see the tail of the stage ring below)
        for (int i = 0; i < halo_bufs; ++i) {
          WAIT(bar_aempty + 8 * ab, aph ^ 1);
          if (++ab == halo_bufs) { ab = 0; aph ^= 1; }
        }
      }
    }
    int lt = 0;
    auto gemm_tiles = [&](auto hint_tag) {
    constexpr bool HINT = decltype(hint_tag)::value;
    for (int tile = tile0; tile < num_tiles && !(TR > 0 && p.b_resident) && mode != 3; tile += tstride, ++lt) {
      PL_TRACE(0, lt, 0);
      int tn_, tm_;
      tile_mn<GROUPED, CTA2>(p, tile, tm_, tn_, cta_rank);
      const int tn = tn_, tm = tm_;
      const int m0 = (tm * kBM) << p.mpair, n0 = tn * BN;  // pairs: one 256-row box
      if (EPI == EPI_I8_RELU_RES) {
        // residual tile first: it is the largest load of a 1x1 expand layer
        WAIT(bar_rempty + 8 * rb, rphase ^ 1);
        PL_TRACE(0, lt, 3);
        {
          mbar_arrive_expect_tx(bar_rfull + 8 * rb, tile_bytes);
          for (int j = 0; j < p.panels; ++j) {
            if (HINT)
              tma_load_2d_hint(smem_res + rb * tile_bytes + j * (kBM * p.PW), &tmap_r, bar_rfull + 8 * rb,
                               n0 + j * p.PW, m0, pol_r);
            else
              tma_load_2d(smem_res + rb * tile_bytes + j * (kBM * p.PW), &tmap_r, bar_rfull + 8 * rb,
                          n0 + j * p.PW, m0);
          }
        }
        if (++rb == p.res_bufs) { rb = 0; rphase ^= 1; }
      }
      int bw = 0, bh = 0, bn = 0;
      if (mode == 1) {
        bn = (int)fdiv((uint32_t)m0, p.dPQ);
        const int rem = m0 - bn * p.PQ;
        const int op = (int)fdiv((uint32_t)rem, p.dQ), oq = rem - op * p.Q;
        bw = oq * p.sw - p.pl_;
        bh = op * p.sh - p.pt;
      }
      int kc = 0, cbk = 0, sx = 0;      // K coordinate of B, channel offset inside the tap, tap column
      uint16_t offw = 0, offh = 0;      // im2col filter offsets of the tap
      for (int kb = 0; kb < num_kb;) {
        // one group of KG ring slots (the last group of a tile may be shorter): the barriers of its first slot
        const int gn = (num_kb - kb) < KG ? (num_kb - kb) : KG;
        WAIT(empty_b, phase ^ 1);
        if (kb == 0) PL_TRACE(0, lt, 1);
        // pairs: own 128 (256) rows of A + own half of every B block, all completing on the leader's barrier
        if (!CTA2 || cta_rank == 0) mbar_arrive_expect_tx(full_b, (CTA2 ? 2u : 1u) * (uint32_t)gn * tx);
        const uint32_t fsig = CTA2 ? full_b + lead_delta : full_b;
        uint32_t da_ = dst_a, db_ = dst_b;
        for (int g = 0; g < gn; ++g, ++kb, da_ += a_bytes, db_ += b_bytes) {
          if (CTA2) {
            if (mode == 1) tma_load_im2col_4d_2sm(da_, &tmap_a, fsig, cbk, bw, bh, bn, offw, offh);
            else tma_load_2d_2sm(da_, &tmap_a, fsig, kc, m0);
            tma_load_2d_2sm(db_, &tmap_b, fsig, kc, n0 + cta_rank * (BN >> 1));
          } else {
            if (TR > 0) {
              // A arrives from the gather warps
            } else if (mode == 1) {
              if (HINT) tma_load_im2col_4d_hint(da_, &tmap_a, full_b, cbk, bw, bh, bn, offw, offh, pol_a);
              else tma_load_im2col_4d(da_, &tmap_a, full_b, cbk, bw, bh, bn, offw, offh);
            } else {
              if (HINT) tma_load_2d_hint(da_, &tmap_a, full_b, kc, m0, pol_a);
              else tma_load_2d(da_, &tmap_a, full_b, kc, m0);
            }
            if (!bres01) tma_load_2d(db_, &tmap_b, full_b, kc, n0);
          }
          kc += BK;
          cbk += BK;
          if (cbk == cspan) {
            cbk = 0;
            offw += dwv;
            if (++sx == S) { sx = 0; offw = 0; offh += dhv; }
          }
        }
        stage += KG; full_b += 8 * KG; empty_b += 8 * KG; dst_a += KG * a_bytes; dst_b += KG * b_bytes;
        if (stage >= stages) {
          stage = 0; phase ^= 1;
          full_b = bar_full; empty_b = bar_empty; dst_a = smem_a; dst_b = smem_b;
        }
      }
      PL_TRACE(0, lt, 2);
    }
    };  // gemm_tiles
    gemm_tiles(std::true_type{});
    if (CTA2) {
      // tail: every stage this CTA filled has been consumed (its multicast commit has landed here) before the
      // CTA can leave -- no arrival may target the barriers of a CTA that has exited
      const int tail_step = (mode == 3 || TR > 0) ? 1 : KG;
      for (int i = 0; i < stages; i += tail_step) {
        WAIT(empty_b, phase ^ 1);
        stage += tail_step; empty_b += 8 * tail_step;
        if (stage >= stages) { stage = 0; phase ^= 1; empty_b = bar_empty; }
      }
    }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // ONE elected lane runs the whole nest.  (Measured alternative: all 32 lanes walking the
    // nest with warp-uniform state in uniform registers and only the MMA under the elected
    // lane is ~30% slower on the 3x3 layers.)  The issue cost per MMA is what bounds the
    // short-K layers, so the k-steps of a block and the taps of a resident 3x3 filter are
    // fully unrolled (issue_kblock / issue_taps).
    // (pairs: only the leader CTA issues; its MMAs run on both SMs)
    if ((!CTA2 || cta_rank == 0) && elect_one()) {
    constexpr bool leader = true;
#define PL_MTRACE(t, e) do { if (leader) PL_TRACE(1, t, e); } while (0)
    const uint32_t idesc = make_idesc_i8((uint32_t)p.BN, CTA2 ? 256u : 128u);
    const uint64_t desc_hi = make_smem_desc_base((uint32_t)p.BK);
    const int ksteps = p.BK / 32, stages = p.stages, num_kb = p.num_kb;
    const uint32_t a_step = p.a_stage_bytes >> 4, b_step = p.b_stage_bytes >> 4;
    const uint32_t a_lo0 = smem_a >> 4, b_lo0 = smem_b >> 4;  // shared addresses < 256 KB: 14 bits
    const bool bres = (TR > 0 || p.mode != 3) && p.b_resident;
    const bool skip_mma = (PL_DBG_FLAGS_V & 2) != 0;
    const int KG = (TR > 0 || p.kgroup < 1) ? 1 : p.kgroup;
    int stage = 0;
    uint32_t phase = 0;
    uint32_t full_b = bar_full, empty_b = bar_empty, a_lo = a_lo0, b_lo = b_lo0;
    int as = 0;
    uint32_t aphase = 0;
    if (bres) WAIT(TR > 0 ? bar_rfull : bar_bres, 0);
    if (TR == 0 && p.mode == 3) {
      const int S = p.S, RS = p.R * p.S, cblocks = p.cblocks;
      const uint32_t tap_step = (uint32_t)p.BK >> 4;                       // next tap column: one pixel
      const uint32_t row_step = (uint32_t)((p.WP - S) * p.BK) >> 4;        // extra step at the end of a filter row
      const uint32_t halo_step = p.a_halo_bytes >> 4;
      const int halo_bufs = p.halo_bufs;
      const bool hres = p.b_resident != 0;
      int ab = 0;
      uint32_t aph = 0;
      if (hres) WAIT(bar_bres, 0);
      int lt = 0;
      if (p.s2d) {
        S2dArgs sa;
        sa.tile0 = tile0; sa.tstride = tstride; sa.num_tiles = num_tiles;
        sa.tmem_base = tmem_base; sa.acc_stride = p.acc_stride; sa.acc_stages = p.acc_stages;
        sa.bar_tempty = bar_tempty; sa.bar_tfull = bar_tfull; sa.bar_afull = bar_afull; sa.bar_aempty = bar_aempty;
        sa.a_lo0 = a_lo0; sa.b_lo0 = b_lo0; sa.halo_step = halo_step; sa.b_step = b_step;
        sa.pitch16 = p.s2d_span_pitch >> 4;
        sa.halo_bufs = halo_bufs; sa.ta = p.s2d_ta; sa.kpr = p.s2d_kpr;
        sa.desc_hi = desc_hi; sa.idesc = idesc; sa.skip_mma = skip_mma ? 1 : 0; sa.rows = p.hk; sa.bn = p.BN;
#ifdef PL_DEBUG_KNOBS
        sa.trace = blockIdx.x == 0 ? p.trace : nullptr;
#endif
        mma_loop_s2d(sa);
      }
      const bool taps3 = hres && RS == 9 && S == 3 && !skip_mma && !(PL_DBG_FLAGS_V & 128);
      if (taps3 && !p.s2d) {
        Taps3Args ta3;
        ta3.tile0 = tile0; ta3.tstride = tstride; ta3.num_tiles = num_tiles;
        ta3.tmem_base = tmem_base; ta3.acc_stride = p.acc_stride; ta3.acc_stages = p.acc_stages;
        ta3.bar_tempty = bar_tempty; ta3.bar_tfull = bar_tfull; ta3.bar_afull = bar_afull; ta3.bar_aempty = bar_aempty;
        ta3.a_lo0 = a_lo0; ta3.b_lo0 = b_lo0; ta3.halo_step = halo_step; ta3.b_step = b_step;
        ta3.row_pitch = (uint32_t)(p.WP * p.BK) >> 4;
        ta3.halo_bufs = halo_bufs; ta3.cblocks = cblocks; ta3.ksteps = ksteps;
        ta3.desc_hi = desc_hi; ta3.idesc = idesc;
#ifdef PL_DEBUG_KNOBS
        ta3.trace = blockIdx.x == 0 ? p.trace : nullptr;
#endif
        mma_loop_taps3<CTA2>(ta3);
      }
      for (int tile = tile0; tile < num_tiles && !p.s2d && !taps3; tile += tstride, ++lt) {
        PL_MTRACE(lt, 0);
        WAIT(bar_tempty + 8 * as, aphase ^ 1);
        PL_MTRACE(lt, 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)as * p.acc_stride;
        if (hres) b_lo = b_lo0;
        for (int cb = 0; cb < cblocks; ++cb) {
          WAIT(bar_afull + 8 * ab, aph);
          if (cb == 0) PL_MTRACE(lt, 2);
          tc_fence_after();
          uint32_t t_lo = a_lo0 + (uint32_t)ab * halo_step;  // descriptor start of tap (0, 0)
          int sx = 0;
          for (int tap = 0; tap < RS; ++tap) {
            if (!hres) {  // streamed filter: one ring stage per tap
              WAIT(full_b, phase);
              tc_fence_after();
            }
            const uint64_t da = desc_hi | (uint64_t)t_lo, db = desc_hi | (uint64_t)b_lo;
            if (leader) {
              if (!skip_mma) issue_kblock_n<CTA2>(ksteps, d_tmem, da, db, idesc, (cb | tap) == 0);
              if (!hres) umma_commit_t<CTA2>(empty_b);
              if (tap == RS - 1) {
                umma_commit_t<CTA2>(bar_aempty + 8 * ab);
                if (cb == cblocks - 1) umma_commit_t<CTA2>(bar_tfull + 8 * as);
              }
            }
            t_lo += tap_step;
            if (++sx == S) { sx = 0; t_lo += row_step; }
            b_lo += b_step;
            if (!hres) {
              ++stage; full_b += 8; empty_b += 8;
              if (stage == stages) { stage = 0; phase ^= 1; full_b = bar_full; empty_b = bar_empty; b_lo = b_lo0; }
            }
          }
          if (++ab == halo_bufs) { ab = 0; aph ^= 1; }
        }
        PL_MTRACE(lt, 4);
        if (++as == p.acc_stages) { as = 0; aphase ^= 1; }
      }
    }
    int lt = 0;
    if (TR > 0 && p.thin_group) {
      // thin mode, resident filter: a ring stage is a whole tile (all k-blocks): one wait, one proxy
      // fence, a straight run of MMAs, two commits per tile
      constexpr int RPK_ = 8 / (TSW > 0 ? TSW : 8), NKB_ = ((TR > 0 ? TR : 1) + RPK_ - 1) / RPK_;
      constexpr uint32_t kb_step = (uint32_t)(kBM * 32) >> 4;
      for (int tile = tile0; tile < num_tiles; tile += tstride, ++lt) {
        PL_MTRACE(lt, 0);
        WAIT(bar_tempty + 8 * as, aphase ^ 1);
        PL_MTRACE(lt, 1);
        WAIT(full_b, phase);
        PL_MTRACE(lt, 2);
        fence_proxy_async();  // cp.async (generic proxy) writes of the gather -> tensor core reads
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)as * p.acc_stride;
        if (!skip_mma) {
#pragma unroll
          for (int kb = 0; kb < NKB_; ++kb)
            umma_i8_t<CTA2>(d_tmem, desc_hi | (uint64_t)(a_lo + (uint32_t)kb * kb_step),
                    desc_hi | (uint64_t)(b_lo0 + (uint32_t)kb * b_step), idesc, (uint32_t)(kb > 0));
        }
        umma_commit_t<CTA2>(empty_b);
        umma_commit_t<CTA2>(bar_tfull + 8 * as);
        ++stage; full_b += 8; empty_b += 8; a_lo += a_step;
        if (stage == stages) { stage = 0; phase ^= 1; full_b = bar_full; empty_b = bar_empty; a_lo = a_lo0; }
        PL_MTRACE(lt, 4);
        if (++as == p.acc_stages) { as = 0; aphase ^= 1; }
      }
    }
    for (int tile = tile0; tile < num_tiles && !(TR == 0 && p.mode == 3) && !(TR > 0 && p.thin_group);
         tile += tstride, ++lt) {
      PL_MTRACE(lt, 0);
      WAIT(bar_tempty + 8 * as, aphase ^ 1);
      PL_MTRACE(lt, 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)as * p.acc_stride;
      if (bres) b_lo = b_lo0;
      for (int kb = 0; kb < num_kb;) {
        const int gn = (num_kb - kb) < KG ? (num_kb - kb) : KG;  // k-blocks of this barrier group
        WAIT(full_b, phase);
        if (kb == 0) PL_MTRACE(lt, 2);
        if (kb == KG) PL_MTRACE(lt, 5);
        if (kb + gn == num_kb) PL_MTRACE(lt, 3);
        if (TR > 0) fence_proxy_async();  // cp.async (generic proxy) writes of the gather -> tensor core reads
        tc_fence_after();
        uint32_t ga = a_lo, gb = b_lo;
        for (int g = 0; g < gn; ++g, ga += a_step, gb += b_step) {
          uint64_t da = desc_hi | (uint64_t)ga;
          const uint64_t db = desc_hi | (uint64_t)gb;
          if (PL_DBG_ASHIFT_V) {  // descriptor-semantics experiment (tools/dbg_rowshift.py)
            const uint32_t start = (ga << 4) + (uint32_t)(PL_DBG_ASHIFT_V * p.BK);
            da = desc_hi | (uint64_t)((start >> 4) & 0x3fff);
            if (PL_DBG_BASEOFF_V) da |= (uint64_t)((start >> 7) & 7u) << 49;
          }
          if (!skip_mma) {
            issue_kblock_n<CTA2>(ksteps, d_tmem, da, db, idesc, kb + g == 0);
            if (p.mpair)  // second 128-row block of the stage, same B, second accumulator
              issue_kblock_n<CTA2>(ksteps, d_tmem + (uint32_t)p.BN, da + (uint64_t)((kBM * p.BK) >> 4), db, idesc, kb + g == 0);
          }
        }
        kb += gn;
        umma_commit_t<CTA2>(empty_b);
        if (kb == num_kb) umma_commit_t<CTA2>(bar_tfull + 8 * as);
        stage += KG; full_b += 8 * KG; empty_b += 8 * KG; a_lo += KG * a_step;
        b_lo += (bres ? gn : KG) * b_step;  // resident filter: consecutive k-blocks; ring: the group's slots
        if (stage >= stages) {
          stage = 0; phase ^= 1;
          full_b = bar_full; empty_b = bar_empty; a_lo = a_lo0;
          if (!bres) b_lo = b_lo0;
        }
      }
      PL_MTRACE(lt, 4);
      if (++as == p.acc_stages) { as = 0; aphase ^= 1; }
    }
    }
    __syncwarp();
  } else if (TR > 0 && warp >= 2 + EW && warp < kStoreWarp) {
    // ===================== thin-input gather (8 warps, two threads per A row) =====================
    constexpr int R_ = TR > 0 ? TR : 1, SW_ = TSW > 0 ? TSW : 4;
    constexpr int RPK = 8 / SW_;                  // filter rows per 32-byte k-block
    constexpr int NKB = (R_ + RPK - 1) / RPK;
    constexpr int MYKB = (NKB + 1) / 2;           // k-blocks per thread: kb = half, half + 2, ...
    const int g = (warp - 2 - EW) * 32 + lane;
    const int prow = g & (kBM - 1), half = g >> 7;
    const uint32_t rowoff = (uint32_t)prow * 32u;
    const uint32_t swx = (rowoff >> 7) & 1u;      // 32-byte swizzle: chunk ^= bit 7 of the offset
    pdl_wait();
    uint32_t base_st = 0, base_ph = 0;  // ring stage / phase of this tile's k-block 0
    int glt = 0;
    for (int tile = tile0; tile < num_tiles; tile += tstride, ++glt) {
      if (g == 0) PL_TRACE(0, glt, 0);
      const int tm = (int)fdiv((uint32_t)(p.rev_last >= 0 ? p.rev_last - tile : tile), p.dNT);
      const uint32_t m = (uint32_t)(tm * kBM + prow);
      const bool mok = m < (uint32_t)p.M;
      const uint32_t n = fdiv(mok ? m : 0u, p.dPQ);
      const uint32_t rem = (mok ? m : 0u) - n * (uint32_t)p.PQ;
      const uint32_t op = fdiv(rem, p.dQ);
      const int oq = (int)(rem - op * (uint32_t)p.Q);
      const int ih0 = (int)op * p.sh - p.pt, iw0 = oq * p.sw - p.pl_ - p.thin_shift;
      const uint32_t* img = p.xw + (size_t)n * (size_t)(p.H * p.W);
      // The window words go global -> shared with cp.async (LDGSTS): nothing is held in registers
      // and nothing waits for the loads, so the gather runs ahead of the tensor core by the whole
      // stage ring.  Zero padding = a clamped (always valid) address with src-size 0, which
      // cp.async zero-fills.  A thread's copies complete its own arrival on the stage's `full`
      // barrier (cp.async.mbarrier.arrive.noinc); the MMA thread orders them against the tensor
      // core's async-proxy reads with a proxy fence after its wait.
      // With a resident filter (p.thin_group) a ring stage holds ALL k-blocks of a tile: one
      // empty-wait and one arrival per thread and tile, one full-wait and one commit per tile in
      // the MMA thread (its try_wait / fence / commit latencies, ~450 cycles per k-block, were
      // the whole cost of this mode when every 4 KB k-block was a stage of its own).
      const uint32_t* cp[SW_];
      uint32_t csz[SW_];
#pragma unroll
      for (int j = 0; j < SW_; ++j) {
        const int iw = iw0 + j, c = min(max(iw, 0), p.W - 1);
        cp[j] = img + (p.thin_pairs ? (c & ~1) : c);  // 8-byte copies need an aligned (if unread) address
        csz[j] = (iw == c && mok) ? 4u : 0u;
      }
      const bool grouped = p.thin_group != 0;
      if (grouped) {
        if (lane == 0) WAIT(bar_empty + 8 * base_st, base_ph ^ 1);
        __syncwarp();
        if (g == 0) PL_TRACE(0, glt, 1);
      }
#pragma unroll
      for (int i = 0; i < MYKB; ++i) {
        const int kb = half + 2 * i;
        if (kb >= NKB) continue;
        // ring position, kept incrementally (a runtime % and / per k-block is expensive here)
        uint32_t st = base_st, ph = base_ph;
        uint32_t dst;
        if (grouped) {
          dst = smem_a + st * p.a_stage_bytes + (uint32_t)kb * (uint32_t)(kBM * 32) + rowoff;
        } else {
          st += (uint32_t)kb;
          if (st >= (uint32_t)p.stages) { st -= (uint32_t)p.stages; ph ^= 1u; }
          // one lane per warp polls; the warp sync hands the result over
          if (lane == 0) WAIT(bar_empty + 8 * st, ph ^ 1);
          __syncwarp();
          if (g == 0 && i == 0) PL_TRACE(0, glt, 1);
          dst = smem_a + st * p.a_stage_bytes + rowoff;
        }
#pragma unroll
        for (int t = 0; t < RPK; ++t) {
          const int r = kb * RPK + t;  // filter row
          const int ih = ih0 + r, hc = min(max(ih, 0), p.H - 1);
          const bool rok = (ih == hc) && r < R_;
          const uint32_t roff = (uint32_t)(hc * p.W);
          if (PL_DBG_FLAGS_V & 32) {
            // ablation: no copies at all (garbage operands)
          } else if (p.thin_pairs) {
            // even window start, even W: a pair of words is inside or outside the row together
#pragma unroll
            for (int j = 0; j < SW_; j += 2) {
              const uint32_t wi = (uint32_t)(t * SW_ + j);
              cp_async_8(dst + (((wi >> 2) ^ swx) << 4) + ((wi & 3u) << 2), cp[j] + roff, rok ? 2u * csz[j] : 0u);
            }
          } else {
#pragma unroll
            for (int j = 0; j < SW_; ++j) {
              const uint32_t wi = (uint32_t)(t * SW_ + j);  // word of the 32-byte row
              cp_async_4(dst + (((wi >> 2) ^ swx) << 4) + ((wi & 3u) << 2), cp[j] + roff, rok ? csz[j] : 0u);
            }
          }
        }
        if (!grouped) cp_async_mbar_arrive_noinc(bar_full + 8 * st);
      }
      if (grouped) cp_async_mbar_arrive_noinc(bar_full + 8 * base_st);
      if (g == 0) PL_TRACE(0, glt, 2);
      base_st += grouped ? 1u : (uint32_t)NKB;  // NKB <= stages (checked on the host)
      if (base_st >= (uint32_t)p.stages) { base_st -= (uint32_t)p.stages; base_ph ^= 1u; }
    }
  } else if (warp == kStoreWarp) {
    // ===================== output store warp =====================
    // One elected lane turns every staged tile into TMA stores.  The epilogue warps never
    // meet at a block barrier: they hand a staging buffer over through `staged` (one arrival
    // per warp) and get it back through `ofree` once the store has read it out.
    if (EPI != EPI_GENERIC && p.tma_out && p.mc_out != nullptr) {
      // multicast variant (mode 0 GEMMs only): the whole warp copies the swizzled staging panels row by row,
      // lane = (row & 3, 16-byte chunk); rows >= M and columns >= Nout are dropped here (no hardware clipping)
      pdl_wait();
      const int nbufs = p.out_bufs;
      int ob = 0;
      uint32_t ophase = 0;
      const uint32_t chunk = (uint32_t)lane & 7u, rsub = (uint32_t)lane >> 3;
      for (int tile = tile0; tile < num_tiles; tile += tstride) {
        int tn_, tm_;
        tile_mn<GROUPED, CTA2>(p, tile, tm_, tn_, cta_rank);
        const int n0 = tn_ * p.BN;
        const uint32_t out_tile = smem_out + (uint32_t)ob * obuf_bytes;
        if (lane == 0) WAIT(bar_staged + 8 * ob, ophase);
        __syncwarp();
        for (int sub = 0; sub <= p.mpair; ++sub) {
          const long long mrow = (long long)((tm_ << p.mpair) + sub) * kBM;
          for (int j = 0; j < p.panels; ++j) {
            const uint32_t panel = out_tile + (uint32_t)sub * tile_bytes + (uint32_t)(j * kBM * p.PW);
            const int col = n0 + j * p.PW + (int)chunk * 16;
            const uint32_t cpr = (uint32_t)p.PW >> 4;  // 16-byte chunks per panel row
            if (chunk < cpr && col < p.Nout) {
              for (uint32_t r = rsub; r < (uint32_t)kBM; r += 4) {
                if (mrow + r >= p.M) break;
                const uint32_t swx = ((r * (uint32_t)p.PW) >> 7) & p.swz_mask;  // as the epilogue staged it
                const uint4 v = ld_shared_v4(panel + r * (uint32_t)p.PW + ((chunk ^ swx) << 4));
                multimem_st_v4(p.mc_out + (mrow + r) * (long long)p.ldy + col, v);
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_ofree + 8 * ob);
        if (++ob == nbufs) { ob = 0; ophase ^= 1; }
      }
    } else if (EPI != EPI_GENERIC && p.tma_out && elect_one()) {
      pdl_wait();
      const bool halo = (TR == 0) && p.mode == 3;
      constexpr bool kInPlace = EPI == EPI_I8_RELU_RES;  // the output sits in the residual ring
      const int nbufs = kInPlace ? p.res_bufs : p.out_bufs;
      // A buffer goes back to the epilogue (producer) only once its TMA store has READ it out; waiting for that
      // after every tile made this one thread the tile clock of the narrow layers (~700-1000 cycles of read-out
      // latency per tile).  Up to `defer` = nbufs - 2 store groups stay in flight instead: the buffer of tile
      // i - defer is released after tile i's stores are committed, and two buffers always remain for the
      // epilogue groups that are writing
      const int defer = p.store_defer;
      int ob = 0, rel = 0, pending = 0;
      uint32_t ophase = 0;
      auto release = [&]() {
        mbar_arrive(kInPlace ? bar_rempty + 8 * rel : bar_ofree + 8 * rel);
        if (++rel == nbufs) rel = 0;
        --pending;
      };
      for (int tile = tile0; tile < num_tiles; tile += tstride) {
        int tn_, tm_;
        tile_mn<GROUPED, CTA2>(p, tile, tm_, tn_, cta_rank);
        const int tn = tn_, tm = tm_;
        const int n0 = tn * p.BN;
        const uint32_t out_tile = (kInPlace ? smem_res : smem_out) + (uint32_t)ob * (kInPlace ? tile_bytes : obuf_bytes);
        WAIT(bar_staged + 8 * ob, ophase);
        if (!(PL_DBG_FLAGS_V & 16)) {
          if (halo) {  // box = PW channels x Q pixels x hk rows of one image; rows >= P are clipped
            const int himg = (int)fdiv((uint32_t)tm, p.dBands);
            const int hp0 = (tm - himg * p.bands) * p.hk;
            for (int j = 0; j < p.panels; ++j)
              tma_store_4d(&tmap_y, out_tile + (uint32_t)(j * kBM * p.PW), n0 + j * p.PW, 0, hp0, himg);
          } else {
            for (int sub = 0; sub <= p.mpair; ++sub) {
              const uint32_t src = out_tile + (uint32_t)sub * tile_bytes;
              const int mrow = ((tm << p.mpair) + sub) * kBM;
              for (int j = 0; j < p.panels; ++j)
                tma_store_2d(&tmap_y, src + (uint32_t)(j * kBM * p.PW), n0 + j * p.PW, mrow);
              for (int pe = 0; pe < p.n_peers; ++pe)
                for (int j = 0; j < p.panels; ++j)
                  tma_store_2d(p.peer_maps + pe, src + (uint32_t)(j * kBM * p.PW), n0 + j * p.PW, mrow);
            }
          }
          bulk_commit_group();
        }
        ++pending;
        if (pending > defer) {
          switch (defer) {
            case 0: bulk_wait_group_read<0>(); break;
            case 1: bulk_wait_group_read<1>(); break;
            case 2: bulk_wait_group_read<2>(); break;
            case 3: bulk_wait_group_read<3>(); break;
            default: bulk_wait_group_read<4>(); break;
          }
          release();
        }
        if (++ob == nbufs) { ob = 0; ophase ^= 1; }
      }
      bulk_wait_group_read<0>();
      while (pending > 0) release();
#ifndef PL_EXP_NO_FINAL_WAIT
      bulk_wait_group<0>();
#endif
    }
    __syncwarp();
  } else {
    // ===================== epilogue (warps 2..17) =====================
    // warp w owns TMEM lanes [32*(w%4), +32) (hardware rule) and the 16-column units
    // cq, cq+CQ, cq+2CQ, ... of the tile (CQ = epilogue warps / 4)
    const int quarter = hw_warp & 3;  // TMEM lane quarter is a property of the hardware warp id
    const int cq = (warp - 2) >> 2;
    const int etid = (warp - 2) * 32 + lane;
    constexpr bool kFast = (EPI != EPI_GENERIC);  // int8 paths: swizzled smem staging + TMA store
    constexpr int CQ = EW / 4;  // warps per TMEM lane quarter
    // accumulator row of this thread -> row of the (dense) output tile.  Halo mode computes
    // on the padded grid: row i is output (i / WP, i % WP) of the band, rows in the pad
    // columns or beyond the band carry garbage and are dropped here
    const int arow = quarter * 32 + lane;
    const bool halo = (TR == 0) && p.mode == 3;
    int orow = 0, jrow = arow;
    bool rvalid = true;
    if (halo) {
      orow = (int)fdiv((uint32_t)arow, p.dWP);
      const int ocol = arow - orow * p.WP;
      rvalid = ocol < p.Q && orow < p.hk;
      jrow = rvalid ? orow * p.Q + ocol : 0;
    }
    const uint32_t rowoff = (uint32_t)(jrow * p.PW);
    const uint32_t swx = (rowoff >> 7) & p.swz_mask;
    int sbuf = 0;  // per-tile scale / bias buffer (very wide outputs only)
    // folded scale / bias live in smem.  Normally the vectors of every n-tile are staged once
    // here; a global load per tile would sit on the epilogue's critical path (the tile rate of
    // short-K layers is 1 / epilogue latency).  Very wide outputs fall back to per-tile staging
    // into two 256-entry buffers, filled during the previous tile.
    auto stage_scale_bias = [&](int t, int buf) {
      if (etid < p.BN) {
        int ttm, ttn;
        tile_mn<GROUPED>(p, t, ttm, ttn);
        const int n = ttn * p.BN + etid;
        const bool ok = n < p.Nout && p.epi.scale != nullptr;
        s_scale[buf * 256 + etid] = ok ? __ldg(p.epi.scale + n) : 0.f;
        s_bias[buf * 256 + etid] = (ok && p.epi.bias) ? __ldg(p.epi.bias + n) : 0.f;
      }
    };
    if (PL_DBG_FLAGS_V & 1024) {
    } else if (p.sb_all) {
      for (int n = etid; n < p.sb_n; n += 32 * EW) {
        const bool ok = n < p.Nout && p.epi.scale != nullptr;
        s_scale[n] = ok ? __ldg(p.epi.scale + n) : 0.f;
        s_bias[n] = (ok && p.epi.bias) ? __ldg(p.epi.bias + n) : 0.f;
      }
    } else if (tile0 < num_tiles) {
      stage_scale_bias(tile0, 0);
    }
    pdl_wait();  // the generic path reads the residual / writes y with plain global accesses
    asm volatile("bar.sync 1, %0;" ::"n"(32 * EW) : "memory");
    // the tile loop is instantiated for the two group layouts in use (compile-time unit stride):
    // one group of all warps, or CQ groups of four warps (one warp per TMEM lane quarter)
    auto epilogue_loop = [&](auto cqn_tag, auto pair_tag) {
    constexpr bool kPair = decltype(pair_tag)::value;  // 256 x BN tile pairs: two accumulators per stage
    constexpr int cqn = decltype(cqn_tag)::value;  // column groups inside one tile group
    constexpr int kSlots = (kPair ? 32 : 16) / CQ;  // unrolled unit slots of one warp (cqn = 1: tiles of <= 16 / CQ units)
    const int units = kPair ? 2 * (p.BN >> 4) : (p.BN >> 4);  // 16-column units of one accumulator stage
    constexpr int G = CQ / cqn;                    // tile-alternating groups
    const int egroup = cq / cqn, cqp = cq - egroup * cqn;
    int lt = egroup;  // local tile counter of this CTA: group g takes tiles g, g + G, ...
    const int trole = (warp == 2 && lane == 0) ? 2 : ((warp == 1 + EW && lane == 0) ? 3 : -1);
#define PL_ETRACE(e) do { if (trole >= 0) PL_TRACE(trole, lt, e); } while (0)
    const int res_bufs_ = p.res_bufs > 0 ? p.res_bufs : 1, out_bufs_ = p.out_bufs > 0 ? p.out_bufs : 1;
    // ring positions of local tile `lt` (the groups interleave on every ring): start at lt = egroup, then
    // advance by G per tile -- kept incrementally, a runtime % and / per ring and tile costs more than
    // the groups save on wide tiles
    // (opaque copies: two registers instead of two constant-bank loads in every unit)
    const uint32_t hi_bits_r = opaque_u32(p.hi_bits);
    const float res_scale_r = __uint_as_float(opaque_u32(__float_as_uint(p.epi.res_scale)));
    // unit slots this warp fills (launch constant): units cqp, cqp + cqn, ... < units
    const int n_my = cqp < units ? (units - cqp + cqn - 1) / cqn : 0;
    // staging offsets of this warp's unit slots (tile-invariant: computed once, not per unit and tile)
    uint32_t soffs[kSlots];
#pragma unroll
    for (int i = 0; i < kSlots; ++i) {
      const uint32_t col_ = (uint32_t)((cqp + cqn * i) * 16);
      soffs[i] = (col_ >> p.pw_shift) * (uint32_t)(kBM * p.PW) + ((((col_ & (uint32_t)(p.PW - 1)) >> 4) ^ swx) << 4);
    }
    int as = egroup % p.acc_stages, rb = egroup % res_bufs_, ob = egroup % out_bufs_;
    uint32_t aphase = (uint32_t)(egroup / p.acc_stages) & 1u, rphase = (uint32_t)(egroup / res_bufs_) & 1u,
             ophase = (uint32_t)(egroup / out_bufs_) & 1u;
    for (int tile = tile0 + egroup * tstride; tile < num_tiles; tile += G * tstride, lt += G) {
      PL_ETRACE(0);
      int tn_, tm_;
      tile_mn<GROUPED, CTA2>(p, tile, tm_, tn_, cta_rank);
      const int tn = tn_, tm = tm_;
      long long row = (long long)(kPair ? 2 * tm : tm) * kBM + arow;  // output pixel index (generic epilogue)
      int himg = 0, hp0 = 0;
      if (halo) {
        himg = (int)fdiv((uint32_t)tm, p.dBands);
        hp0 = (tm - himg * p.bands) * p.hk;
        const int prow = hp0 + orow;
        // an out-of-band row is parked beyond M so that the generic path skips it
        row = (rvalid && prow < p.P) ? ((long long)himg * p.PQ + (long long)prow * p.Q + (jrow - orow * p.Q))
                                            : (long long)p.M;
      }
      const int n0 = tn * p.BN;
      const uint32_t sb0 = p.sb_all ? (uint32_t)n0 : (uint32_t)(sbuf * 256);
      // one lane per warp polls (512 threads polling one mbarrier serialise); the warp-level
      // sync hands the acquired state to the other lanes
      // (two lanes poll side by side: a wait costs ~70-250 cycles even when its phase is complete)
      if (lane == 0) {
        WAIT(bar_tfull + 8 * as, aphase);
      } else if (lane == 1) {
        if (EPI == EPI_I8_RELU_RES) WAIT(bar_rfull + 8 * rb, rphase);
        else if (kFast) WAIT(bar_ofree + 8 * ob, ophase ^ 1);
      }
      __syncwarp();
      PL_ETRACE(1);
      tc_fence_after();
      // (per-tile bases; every unit slot is a compile-time offset from them)
      const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)as * p.acc_stride +
                              (uint32_t)(cqp * 16);
      // (opaque: otherwise the compiler re-derives both addresses in every unit instead of keeping two registers)
      const uint32_t sc_a = opaque_u32(smem_u32(&s_scale[sb0 + (uint32_t)(cqp * 16)]));
      const uint32_t bi_a = opaque_u32(smem_u32(&s_bias[sb0 + (uint32_t)(cqp * 16)]));
      const uint32_t res_base = smem_res + (uint32_t)rb * tile_bytes + rowoff;
      const uint32_t out_tile = smem_out + (uint32_t)ob * obuf_bytes;
      // fused residual: each 16-byte chunk of the output replaces the residual chunk it was computed from
      const uint32_t out_base = EPI == EPI_I8_RELU_RES ? res_base : out_tile + rowoff;
      // TMEM reads are double buffered: unit i+1 is in flight while unit i runs through the ALUs
      uint32_t accb[2][16];
      const bool skip_epi = (PL_DBG_FLAGS_V & 1) != 0;
      if (n_my > 0 && !skip_epi) tmem_ld_32x32b_x16(t_addr, accb[0]);
#pragma unroll
      for (int i = 0; i < kSlots; ++i) {
        const int u = cqp + cqn * i;
        if (i < n_my && !skip_epi) {  // warp-uniform
          tmem_ld_wait();
          if (i + 1 < kSlots && i + 1 < n_my)
            tmem_ld_32x32b_x16(t_addr + (uint32_t)(cqn * (i + 1) * 16), accb[(i + 1) & 1]);
          uint32_t(&acc)[16] = accb[i & 1];
          const uint32_t colt = (uint32_t)(u * 16);                       // TMEM column of the unit
          const uint32_t sub = (kPair && colt >= (uint32_t)p.BN) ? 1u : 0u;  // 128-row block of a pair
          const uint32_t col = colt - sub * (uint32_t)p.BN;              // output column inside the tile
          const uint32_t sbase = sub * p.pair_sbase;
          if (kFast) {
            // position of this 16-byte chunk inside the swizzled staging panels; rows >= M
            // and columns >= Nout are computed on zero-filled operands and clipped by the store.  (Padded-grid rows
            // without an output pixel run through the arithmetic too and only skip the store: a divergence region
            // around the body cost every unit three control instructions.)
            const uint32_t soff = kPair ? (col >> p.pw_shift) * (uint32_t)(kBM * p.PW) +
                                              ((((col & (uint32_t)(p.PW - 1)) >> 4) ^ swx) << 4)
                                        : soffs[i];
            uint4 rw = make_uint4(0, 0, 0, 0);
            if (EPI == EPI_I8_RELU_RES) rw = ld_shared_v4(res_base + soff);  // (pairs never carry a residual)
            const uint32_t rws[4] = {rw.x, rw.y, rw.z, rw.w};
            uint32_t w[4];
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float4 sc = kPair ? *reinterpret_cast<const float4*>(&s_scale[sb0 + col + 4 * q4])
                                      : ld_shared_f4(sc_a + (uint32_t)((cqn * i * 16 + 4 * q4) * 4));
              const float4 bi = kPair ? *reinterpret_cast<const float4*>(&s_bias[sb0 + col + 4 * q4])
                                      : ld_shared_f4(bi_a + (uint32_t)((cqn * i * 16 + 4 * q4) * 4));
              if (EPI == EPI_I8_RELU)
                w[q4] = quad_i8_relu<false>(acc + 4 * q4, sc, bi, hi_bits_r, 0u, 0.f);
              else if (EPI == EPI_I8_RELU_RES)
                w[q4] = quad_i8_relu<true>(acc + 4 * q4, sc, bi, hi_bits_r, rws[q4], res_scale_r);
              else
                w[q4] = quad_i8_signed(acc + 4 * q4, sc, bi);
            }
            st_shared_v4_if(out_base + sbase + soff, w[0], w[1], w[2], w[3], rvalid);
          } else if (p.f32_fast && row + (long long)(sub * kBM) < p.M && n0 + (int)col + 16 <= p.Nout) {
            float* out = reinterpret_cast<float*>(p.y) + (row + (long long)(sub * kBM)) * p.ldy + n0 + (int)col;
            const bool relu = p.epi.act == PL_ACT_RELU;
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float4 sc = *reinterpret_cast<const float4*>(&s_scale[sb0 + col + 4 * q4]);
              const float4 bi = *reinterpret_cast<const float4*>(&s_bias[sb0 + col + 4 * q4]);
              float4 v;
              v.x = __fmaf_rn(__int2float_rn((int)acc[4 * q4 + 0]), sc.x, bi.x);
              v.y = __fmaf_rn(__int2float_rn((int)acc[4 * q4 + 1]), sc.y, bi.y);
              v.z = __fmaf_rn(__int2float_rn((int)acc[4 * q4 + 2]), sc.z, bi.z);
              v.w = __fmaf_rn(__int2float_rn((int)acc[4 * q4 + 3]), sc.w, bi.w);
              if (relu) {
                v.x = v.x > 0.f ? v.x : 0.f;
                v.y = v.y > 0.f ? v.y : 0.f;
                v.z = v.z > 0.f ? v.z : 0.f;
                v.w = v.w > 0.f ? v.w : 0.f;
              }
              *reinterpret_cast<float4*>(out + 4 * q4) = v;
            }
          } else if (row + (long long)(sub * kBM) < p.M && n0 + (int)col < p.Nout) {
            store_group16(p, acc, row + (long long)(sub * kBM), n0 + (int)col, &s_scale[sb0 + col], &s_bias[sb0 + col]);
          }
        }
      }
      // the warp's TMEM reads (tcgen05.wait::ld), residual reads and staging writes are complete:
      // one lane per warp releases the accumulator and the residual buffer and hands the
      // staging buffer to the store warp.  No block-level barrier: warps drift apart freely,
      // a fast warp is already reading the next accumulator while a slow one finishes.
      PL_ETRACE(2);
      tc_fence_before();
      if (kFast && !(PL_DBG_FLAGS_V & 64)) fence_proxy_async();  // this thread's st.shared -> visible to the TMA engine
      __syncwarp();
      if (lane == 0) {
        if (CTA2) mbar_arrive_cluster(bar_tempty + 8 * as + lead_delta);  // the leader's MMA thread waits for both CTAs
        else mbar_arrive(bar_tempty + 8 * as);
        if (EPI == EPI_I8_RELU_RES) mbar_arrive(bar_staged + 8 * rb);
        else if (kFast) mbar_arrive(bar_staged + 8 * ob);
      }
      PL_ETRACE(3);
      as += G;
      while (as >= p.acc_stages) { as -= p.acc_stages; aphase ^= 1u; }
      if (EPI == EPI_I8_RELU_RES) {
        rb += G;
        while (rb >= res_bufs_) { rb -= res_bufs_; rphase ^= 1u; }
      } else if (kFast) {
        ob += G;
        while (ob >= out_bufs_) { ob -= out_bufs_; ophase ^= 1u; }
      }
      if (!p.sb_all) {
        // very wide outputs: next tile's scale / bias into the other buffer, published by a
        // barrier of the epilogue warps (this mode keeps the warps in step)
        sbuf ^= 1;
        if (tile + tstride < num_tiles) stage_scale_bias(tile + tstride, sbuf);
        asm volatile("bar.sync 2, %0;" ::"n"(32 * EW) : "memory");
      }
      PL_ETRACE(6);
    }
    };  // epilogue_loop
    bool pair_done = false;
    if constexpr (TR == 0 && (EPI == EPI_I8_RELU || EPI == EPI_I8_SIGNED || EPI == EPI_GENERIC)) {
      if (p.mpair) {
        epilogue_loop(std::integral_constant<int, CQ>{}, std::true_type{});
        pair_done = true;
      }
    }
    if (pair_done) {
    } else if (p.epi_groups > 1) epilogue_loop(std::integral_constant<int, 1>{}, std::false_type{});
    else epilogue_loop(std::integral_constant<int, CQ>{}, std::false_type{});
  }

  tc_fence_before();
  if (CTA2) cluster_sync_all();  // neither CTA's smem / TMEM / barriers go away while the other still uses them
  else __syncthreads();
  if (warp == 1 && !(PL_DBG_FLAGS_V & 512)) {
    tc_fence_after();
    if (CTA2) tmem_dealloc_2sm(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// --------------------------------------------------------------------------
// tensor-pipe micro-benchmark: back-to-back MMAs on resident smem operands
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) i8_mma_peak_kernel(int iters, int bn, int32_t* out, int row_bytes,
                                                             int a_shift_rows) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // (a_shift_rows: the A descriptor starts that many rows into the tile, as the halo-mode taps do)
  const uint32_t sa = base + (uint32_t)(a_shift_rows * row_bytes), sb = base + 160 * 128;
  // deterministic small operands so the checksum is reproducible
  for (uint32_t i = threadIdx.x; i < (160u + (uint32_t)bn) * 128u / 4u; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x01010101u));
  fence_proxy_async();
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0 && elect_one()) {
    const uint32_t idesc = make_idesc_i8((uint32_t)bn);
    // row_bytes selects the operand layout (swizzle mode); one iteration = 8 MMAs cycling over
    // the row_bytes / 32 K-steps of the tile, issued from the uniform datapath
    const uint64_t hi = make_smem_desc_base((uint32_t)row_bytes);
    const uint64_t da = hi | (uint64_t)((sa >> 4) & 0x3fff), db = hi | (uint64_t)((sb >> 4) & 0x3fff);
    const uint32_t kmask = (uint32_t)(row_bytes / 32 - 1);
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tmem + (uint32_t)((i & 1) * bn);
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint64_t ko = (uint64_t)(2u * (k & kmask));
        umma_i8(d, da + ko, db + ko, idesc, (uint32_t)(i > 1 || k > 0));
      }
    }
    umma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  uint32_t r[16];
  tmem_ld_32x32b_x16(tmem + ((uint32_t)(warp * 32) << 16), r);
  tmem_ld_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (int32_t)r[0];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// the same probe for CTA pairs: clusters of two CTAs, the leader issues M = 256 MMAs whose B operand (bn rows) is
// split over the two CTAs' shared memories; counts as 2 x (128 x bn x 32) per instruction
__global__ void __launch_bounds__(128, 1) i8_mma_peak_pair_kernel(int iters, int bn, int32_t* out, int row_bytes) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sa = base, sb = base + 128 * 128;
  for (uint32_t i = threadIdx.x; i < (128u + (uint32_t)bn / 2u) * 128u / 4u; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x01010101u));
  fence_proxy_async();
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc_2sm(smem_u32(&tmem_slot), 512);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (rank == 0 && warp == 0 && elect_one()) {
    const uint32_t idesc = make_idesc_i8((uint32_t)bn, 256u);
    const uint64_t hi = make_smem_desc_base((uint32_t)row_bytes);
    const uint64_t da = hi | (uint64_t)((sa >> 4) & 0x3fff), db = hi | (uint64_t)((sb >> 4) & 0x3fff);
    const uint32_t kmask = (uint32_t)(row_bytes / 32 - 1);
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tmem + (uint32_t)((i & 1) * bn);
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint64_t ko = (uint64_t)(2u * (k & kmask));
        umma_i8_2sm(d, da + ko, db + ko, idesc, (uint32_t)(i > 1 || k > 0));
      }
    }
    umma_commit_2sm(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  uint32_t r[16];
  tmem_ld_32x32b_x16(tmem + ((uint32_t)(warp * 32) << 16), r);
  tmem_ld_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (int32_t)r[0];
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem, 512);
  }
}

// --------------------------------------------------------------------------
// filter packing: OHWI [K,R,S,C] -> [K][R*S][Cb*BK], zero padded
// --------------------------------------------------------------------------
__global__ void pack_conv_weight_kernel(const int8_t* __restrict__ w, int8_t* __restrict__ out,
                                        int K, int RS, int C, int cpad) {
  const long long total = (long long)K * RS * cpad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cpad);
    const long long kt = i / cpad;  // k * RS + tap
    out[i] = c < C ? w[kt * C + c] : (int8_t)0;
  }
}

// thin-input filter packing: OHWI [K,R,S,C<=4] -> [K][NKB][32 bytes]; k-block kb holds
// 32/(4*SW) filter rows of SW words, word j of a row = the 4 channel weights of tap s = j
__global__ void pack_thin_weight_kernel(const int8_t* __restrict__ w, int8_t* __restrict__ out, int K, int R,
                                        int S, int C, int SW, int row_bytes, int shift) {
  const long long total = (long long)K * row_bytes;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / row_bytes), b = (int)(i % row_bytes);
    const int word = b >> 2, c = b & 3;
    const int r = word / SW, s_ = word % SW - shift;
    out[i] = (r < R && s_ >= 0 && s_ < S && c < C) ? w[(((long long)k * R + r) * S + s_) * C + c] : (int8_t)0;
  }
}

// space-to-depth stem filter packing: OHWI [K,R,S,C<=4] -> [K][TA*TB/2][32 bytes]; k-block
// (a, b/2) holds s2d taps b, b+1 of tap row a; byte (b&1)*16 + (dy*2+dx)*4 + c of a k-block is the
// weight of original tap (r, s) = (2a + dy - shift, 2b + dx - shift), zero outside the window
__global__ void pack_s2d_weight_kernel(const int8_t* __restrict__ w, int8_t* __restrict__ out, int K, int R,
                                       int S, int C, int TA, int TB, int shift) {
  const int row_bytes = TA * TB * 16;
  const long long total = (long long)K * row_bytes;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(i / row_bytes), byte = (int)(i % row_bytes);
    const int pix = byte >> 4, sub = byte & 15;
    const int a = pix / TB, b = pix % TB;
    const int dy = sub >> 3, dx = (sub >> 2) & 1, c = sub & 3;
    const int r = 2 * a + dy - shift, s_ = 2 * b + dx - shift;
    out[i] = (r >= 0 && r < R && s_ >= 0 && s_ < S && c < C) ? w[(((long long)k * R + r) * S + s_) * C + c] : (int8_t)0;
  }
}

// --------------------------------------------------------------------------
// host: tensor maps
// --------------------------------------------------------------------------
// tile order of the next launches of this thread (pl_set_tile_order); PL_TILE_ZIGZAG=1 (experiments) alternates per launch
thread_local int g_tile_order = 0;
thread_local unsigned g_zigzag_count = 0;
bool next_tile_order_reversed() {
  static const bool zigzag = getenv("PL_TILE_ZIGZAG") != nullptr && atoi(getenv("PL_TILE_ZIGZAG")) != 0;
  if (zigzag) return (g_zigzag_count++ & 1u) != 0;
  return g_tile_order != 0;
}
extern "C" void pl_set_tile_order(int reverse) { pl::g_tile_order = reverse ? 1 : 0; }

namespace {

struct DriverFns {
  PFN_cuTensorMapEncodeTiled_v12000 tiled = nullptr;
  PFN_cuTensorMapEncodeIm2col_v12000 im2col = nullptr;
  int sm_count = 0;
  int cc_major = 0;
  int driver_version = 0;
  bool ok = false;
};

DriverFns& driver() {
  // entry points are process-wide; SM count / compute capability belong to the CURRENT device
  static DriverFns per_dev[64];
  static bool ready[64] = {};
  static std::mutex mu;
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  std::lock_guard<std::mutex> lock(mu);
  if (!ready[dev]) {
    DriverFns f;
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      f.tiled = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      f.im2col = reinterpret_cast<PFN_cuTensorMapEncodeIm2col_v12000>(fn);
    cudaDeviceGetAttribute(&f.sm_count, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&f.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDriverGetVersion(&f.driver_version);
    f.ok = f.tiled && f.im2col;
    per_dev[dev] = f;
    ready[dev] = true;
  }
  return per_dev[dev];
}

CUtensorMapSwizzle swizzle_for(int bk) {
  return bk == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                   : (bk == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                               : (bk == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE));
}

// 2-D K-major byte matrix [rows, cols] with row pitch ld; box = [box_rows, bk]
int encode_2d(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint64_t ld, int bk,
              int box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld};
  cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = driver().tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims,
                              strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bk),
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PL_OK : PL_ERR_DRIVER;
}

// 4-D tiled map over an NHWC byte tensor {C, W, H, N} (pixel pitch ld); box {bc, bw, bh, 1};
// out-of-bounds box elements read as zero / are not written
int encode_4d(CUtensorMap* m, const void* base, int C, int W, int H, int N, int ld, int bc, int bw, int bh) {
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)ld, (cuuint64_t)ld * W, (cuuint64_t)ld * W * H};
  cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = driver().tiled(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bc),
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PL_OK : PL_ERR_DRIVER;
}

int encode_im2col(CUtensorMap* m, const void* x, const pl_conv_desc* d, int bk) {
  cuuint64_t dims[4] = {(cuuint64_t)d->C, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
  cuuint64_t strides[3] = {(cuuint64_t)d->ldx, (cuuint64_t)d->ldx * d->W,
                           (cuuint64_t)d->ldx * d->W * d->H};
  int lower[2] = {-d->pad_l, -d->pad_t};
  int upper[2] = {d->pad_r - (d->S - 1) * d->dil_w, d->pad_b - (d->R - 1) * d->dil_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)d->stride_w, (cuuint32_t)d->stride_h, 1};
  CUresult r = driver().im2col(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(x), dims,
                               strides, lower, upper, (cuuint32_t)bk, (cuuint32_t)kBM, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bk),
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return PL_ERR_DRIVER;
  // Same guard CUTLASS applies (cute/atom/copy_traits_sm90_im2col.hpp): drivers
  // <= 13.1 mis-encode im2col maps of tensors smaller than 128 KiB.
  if (driver().driver_version <= 13010 &&
      (uint64_t)d->ldx * d->W * d->H * d->N < 131072ull)
    reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
  return PL_OK;
}

uint32_t pow2_cols(int n) {
  uint32_t c = 32;
  while ((int)c < n) c <<= 1;
  return c;
}

int env_int(const char* name, int dflt);
static inline bool trace_enabled() { return env_int("PL_DBG_TRACE_ON", 0) != 0; }

int pick_bn(int Nout, int m_tiles, int sms) {
  // widest tile that does not over-cover Nout, narrowed until the grid fills the SMs
  const int n16 = (Nout + 15) / 16 * 16;
  const int forced = env_int("PL_FORCE_BN", 0);  // tests: wide tiles on problems small enough for the oracle
  if (forced >= 16 && forced <= 256 && (forced & (forced - 1)) == 0) return forced;
  int cap = 256;
  while (cap > 16 && cap / 2 >= n16) cap >>= 1;
  int bn = cap;
  // (stop when the grid already covers most SMs and the next halving would only add a second, partial wave:
  // a skinny 4096 x 4096 GEMM is latency bound per CTA, 128 CTAs of N = 32 finish in half the time of 256 of N = 16)
  const int fill = env_int("PL_BN_FILL_PCT", 60);
  while (bn > 16) {
    const int tiles = m_tiles * ((Nout + bn - 1) / bn);
    if (tiles >= sms) break;
    const int next = m_tiles * ((Nout + bn / 2 - 1) / (bn / 2));
    if (next > sms && tiles * 100 >= sms * fill) break;
    bn >>= 1;
  }
  return bn;
}

template <int EPI, int TR = 0, int TSW = 0, bool GROUPED = false, bool CTA2 = false>
int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& ty, const CUtensorMap& tr,
               const IgemmParams& p, int grid, size_t smem, pl_stream_t stream) {
  // cudaFuncSetAttribute is per DEVICE: one flag per device ordinal (a process may drive several GPUs)
  static bool configured_dev[64] = {};
  int dev_ = 0;
  cudaGetDevice(&dev_);
  bool& configured = configured_dev[dev_ & 63];
  if (!configured) {
    // 227 KB is the per-CTA limit for static + dynamic shared memory together
    PL_CUDA_OK(cudaFuncSetAttribute(igemm_i8_kernel<EPI, TR, TSW, GROUPED, CTA2>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(221 * 1024)));
    configured = true;
  }
  if (CTA2) {
    if (trace_enabled()) return PL_ERR_UNSUPPORTED;
    launch_pdl_cluster(igemm_i8_kernel<EPI, TR, TSW, GROUPED, CTA2>, dim3(grid), dim3(kThreads), smem,
                       (cudaStream_t)stream, 2, ta, tb, ty, tr, p);
    ++g_launches;
    return launch_status();
  }
  const int threads = TR > 0 ? kThinThreads : kThreads;
  const char* trace_path = trace_enabled() ? getenv("PL_DBG_TRACE") : nullptr;
  if (trace_path && *trace_path) {
    // debugging aid: synchronous launch, CTA 0's timeline written as text (role tile e0..e7)
    static long long* dbuf = nullptr;
    const size_t n = 4 * 64 * 8;
    if (!dbuf) PL_CUDA_OK(cudaMalloc(&dbuf, n * sizeof(long long)));
    PL_CUDA_OK(cudaMemsetAsync(dbuf, 0, n * sizeof(long long), (cudaStream_t)stream));
    IgemmParams q = p;
    q.trace = dbuf;
    igemm_i8_kernel<EPI, TR, TSW, GROUPED><<<grid, threads, smem, (cudaStream_t)stream>>>(ta, tb, ty, tr, q);
    ++g_launches;
    PL_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    std::vector<long long> h(n);
    PL_CUDA_OK(cudaMemcpy(h.data(), dbuf, n * sizeof(long long), cudaMemcpyDeviceToHost));
    if (FILE* fp = fopen(trace_path, "w")) {
      fprintf(fp, "# mode %d BN %d BK %d num_kb %d tiles %d grid %d stages %d out_bufs %d res_bufs %d halo_bufs %d b_resident %d\n",
              p.mode, p.BN, p.BK, p.num_kb, p.m_tiles * p.n_tiles, grid, p.stages, p.out_bufs, p.res_bufs, p.halo_bufs, p.b_resident);
      for (int r = 0; r < 4; ++r)
        for (int t = 0; t < 64; ++t) {
          fprintf(fp, "%d %d", r, t);
          for (int e = 0; e < 8; ++e) fprintf(fp, " %lld", h[(r * 64 + t) * 8 + e]);
          fprintf(fp, "\n");
        }
      fclose(fp);
    }
    return launch_status();
  }
  launch_pdl(igemm_i8_kernel<EPI, TR, TSW, GROUPED>, dim3(grid), dim3(threads), smem, (cudaStream_t)stream, ta, tb, ty, tr, p);
  ++g_launches;
  return launch_status();
}

// Developer knobs (PL_*): read from the environment ONCE per knob and process, not ~20 getenv calls per launch
// (the kernel classes run eagerly, layer by layer).  pl_refresh_knobs() drops the cache -- tests that flip a knob
// between two launches call it.  Keyed by the address of the name literal.
struct KnobCache {
  static constexpr int kMax = 64;
  const char* name[kMax];
  int val[kMax];
  bool set[kMax];
  int n = 0;
  std::mutex mu;
};
KnobCache& knobs() {
  static KnobCache k;
  return k;
}
int env_int(const char* name, int dflt) {
  KnobCache& k = knobs();
  std::lock_guard<std::mutex> lock(k.mu);
  for (int i = 0; i < k.n; ++i)
    if (k.name[i] == name) return k.set[i] ? k.val[i] : dflt;
  const char* v = getenv(name);
  const bool set = v && *v;
  const int val = set ? atoi(v) : 0;
  if (k.n < KnobCache::kMax) {
    k.name[k.n] = name;
    k.val[k.n] = val;
    k.set[k.n] = set;
    ++k.n;
  }
  return set ? val : dflt;
}
bool tma_epilogue_enabled() { return env_int("PL_IGEMM_NO_TMA_EPI", 0) != 1; }

// the B operand as the caller encoded it (K-major [rows, cols] bytes, pitch ld): pair launches re-encode its
// tensor map with a box of BN / 2 rows
struct BSpec {
  const void* base;
  uint64_t cols, rows, ld;
};

// clusters of two CTAs that can be resident at once (normally SMs / 2), per device
int max_pair_clusters();

extern thread_local int g_last_plan[10];

// CTA pairs pay where the operand stream from L2 bounds the launch: long reductions
// whose filter slice cannot stay resident in one CTA's shared memory.  Measured on B200 (tools/gpu_cta2_ab.sh):
// +4..10 % on the ResNet-50 layers with K_red >= 1024 (14x14 / 7x7 and the stride-2 3x3s), 57 -> 64 %, 63 -> 70 %
// and 60 -> 71 % of the tensor peak on 4096^3, 8192^3 and 8192 x 4096 x 8192; launches that keep their filter
// resident lose 20-50 % when they give that up, K_red = 512 is a wash.
bool want_cta_pairs(const IgemmParams& p) {
  const int knob = env_int("PL_CTA2", 1);  // 0 never, 1 auto, 2 whenever the launch qualifies
  const bool legal = knob > 0 && (p.mode == 0 || p.mode == 1 || (p.mode == 3 && !p.s2d)) && p.BN >= 64 &&
                     (p.BN % 32) == 0 && p.m_tiles >= 2 && !trace_enabled() && max_pair_clusters() > 0;
  if (!legal) return false;
  if (knob >= 2) return true;
  const long long kred = (long long)p.num_kb * p.BK;
  // would the single-CTA launch keep its filter resident?  (same tests as below, on the single-CTA geometry)
  const long long fbytes = (long long)p.num_kb * p.b_stage_bytes;
  const int sms = driver().sm_count;
  const int tiles = p.m_tiles * p.n_tiles;
  const int n_cover = p.n_tiles * p.BN;
  const long long budget = 219 * 1024 - ((n_cover <= 2048 ? n_cover : 512) * 8 + 1023) / 1024 * 1024;
  const long long staging = (long long)kBM * p.BN * (p.epi.residual ? 2 : 1);
  if (p.mode == 3 && p.n_tiles == 1) {
    // halo launches with the whole filter in shared memory: a pair keeps HALF of it per CTA.  At N <= 128 the operand
    // reads of the MMAs alone need ~122 B/cycle of the SM's 128 B/cycle shared-memory bandwidth (A 4 KB + B N x 32 B
    // per 67-cycle MMA); in a pair a CTA serves its A and half of B (90 B/cycle), which leaves room for the TMA
    // writes and the epilogue's staging traffic (profiles/r02_ncu_layers.md: pipe 95 % occupied, 49 % delivered)
    const long long half = (long long)p.num_kb * (((long long)(p.BN / 2) * p.BK + 1023) / 1024 * 1024);
    const int hres_knob = env_int("PL_CTA2_HRES", 0);  // 0 off (default: measured neutral), 1 BN >= 128, 2 BN >= 64
    if (2ll * p.a_halo_bytes + half + staging <= budget && hres_knob > 0 && p.BN >= (hres_knob >= 2 ? 64 : 128))
      return true;
    if (2ll * p.a_halo_bytes + fbytes + staging <= budget) return false;
  }
  if (kred < env_int("PL_CTA2_MIN_K", 1024) || p.BN < env_int("PL_CTA2_MIN_BN", 128)) return false;
  if (p.mode != 3 && !p.mpair && tiles >= 2 * sms && p.n_tiles <= 16 && fbytes <= 64 * 1024 &&
      p.epi.out_dtype == PL_OUT_INT8)
    return false;
  return true;
}

int launch_igemm(const CUtensorMap& ta, const CUtensorMap& tb_in, IgemmParams& p, pl_stream_t stream,
                 const BSpec* bs = nullptr) {
  CUtensorMap tb = tb_in;
  p.cta2 = 0;
  {
    if (bs && want_cta_pairs(p)) {
      const uint32_t half = (uint32_t)((p.BN / 2) * p.BK);
      if (encode_2d(&tb, bs->base, bs->cols, bs->rows, bs->ld, p.BK, p.BN / 2) == PL_OK) {
        p.cta2 = 1;
        p.m_tiles = (p.m_tiles + 1) / 2;
        p.b_stage_bytes = (half + 1023u) / 1024u * 1024u;
        p.tx_bytes = p.mode == 3 ? half : p.a_stage_bytes + half;  // per CTA; the leader's barrier expects both
      } else {
        tb = tb_in;
      }
    }
  }
  const int tiles = p.m_tiles * p.n_tiles;
  p.dNT = make_fastdiv((uint32_t)p.n_tiles);
  if (p.mpair && p.pair_sbase == 0) p.pair_sbase = (uint32_t)(kBM * p.BN);
  p.rev_last = next_tile_order_reversed() ? tiles - 1 : -1;
  {
    int knob = env_int("PL_L2_HINTS", 31);
    {
      // only where the launch's tensors are large next to the 126 MB L2 (the hinted loop costs thin-tile launches ~1 %)
      const long long in_b = (long long)p.M * p.cblocks * p.BK, out_b = (long long)p.M * p.Nout;
      const long long all_b = in_b + out_b * (p.epi.residual ? 2 : 1);
      if (all_b < (long long)env_int("PL_L2_HINT_MIN_MB", 48) << 20) knob = 0;
    }
    // activations are read once per layer (a few times within microseconds where tiles overlap): evict_first keeps
    // them from displacing the output tensor, which the next layer reads (pairs keep the default: their layers'
    // tensors fit the L2 anyway).  1 fused residual, 2 A rows of plain GEMMs / 1x1 convs, 4 halo tiles, 8 im2col
    // (large GEMMs re-read A once per N tile column over a long time and live on those L2 hits: no hint there)
    p.l2_hints = (knob & 1) | ((knob & 2) && p.mode == 0 && !p.cta2 && p.n_tiles <= 8 ? 2 : 0) | ((knob & 4) && p.mode == 3 && !p.cta2 ? 4 : 0) |
                 ((knob & 8) && p.mode == 1 && !p.cta2 ? 8 : 0);
  }
  int grid = tiles < driver().sm_count ? tiles : driver().sm_count;
  if (p.cta2) {
    const int clusters = tiles < max_pair_clusters() ? tiles : max_pair_clusters();
    grid = 2 * clusters;
  }
  // epilogue specialisation (ragged edges inside a fast-path launch use the generic code)
  const EpiParams& e = p.epi;
  int epi = EPI_GENERIC;
  if (e.out_dtype == PL_OUT_INT8 && e.round_mode == PL_ROUND_ARM && p.vec_ok && e.scale &&
      tma_epilogue_enabled()) {
    if (e.act == PL_ACT_RELU || e.act == PL_ACT_RELU6) {
      float hi = 127.f;
      if (e.act == PL_ACT_RELU6 && e.act_arg < hi) hi = e.act_arg;
      if (hi > 0.f) {
        union { float f; uint32_t u; } cv;
        cv.f = hi;
        p.hi_bits = cv.u;
        epi = e.residual ? (p.res_vec_ok ? EPI_I8_RELU_RES : EPI_GENERIC) : EPI_I8_RELU;
      }
    } else if (e.act == PL_ACT_NONE && !e.residual) {
      epi = EPI_I8_SIGNED;
    }
  }
  // (tile pairs storing straight from registers through the generic epilogue, to trade the 64 KB of staging
  // for two more mainloop stages, measured 2.5x slower at 4096^3: the exposed epilogue dominates)
  p.f32_fast = (epi == EPI_GENERIC && e.out_dtype == PL_OUT_FP32 && e.round_mode == PL_ROUND_ARM && p.vec_ok &&
                e.scale && !e.residual && (e.act == PL_ACT_NONE || e.act == PL_ACT_RELU) &&
                env_int("PL_NO_F32_FAST", 0) == 0)
                   ? 1
                   : 0;
  // int8 fast paths leave through swizzled smem panels + TMA stores (and take the
  // residual in through TMA loads): vec_ok already guarantees the 16-byte rules
  CUtensorMap ty{}, tr{};
  p.PW = p.BN < 128 ? p.BN : 128;
  p.panels = p.BN / p.PW;
  p.pw_shift = 0;
  while ((1 << p.pw_shift) < p.PW) ++p.pw_shift;
  p.swz_mask = p.PW == 128 ? 7u : (p.PW == 64 ? 3u : (p.PW == 32 ? 1u : 0u));
  p.tma_out = (epi != EPI_GENERIC);
  p.tma_res = p.tma_out && epi == EPI_I8_RELU_RES;
  const int n_cover = p.n_tiles * p.BN;
  p.sb_all = n_cover <= 2048 ? 1 : 0;
  p.sb_n = p.sb_all ? n_cover : 512;
  const int sb_bytes = (p.sb_n * 8 + 1023) / 1024 * 1024;
  const int budget = 219 * 1024 - sb_bytes;  // static smem is now only the barriers
  const int tile = (kBM * p.BN) << p.mpair;  // staging buffer (both blocks of a pair)
  const int stage_bytes = (int)(p.a_stage_bytes + p.b_stage_bytes);
  p.out_bufs = 0;
  p.res_bufs = 2;
  int extra = 0;
  // halo mode with a single N tile keeps the whole packed filter in smem when it fits
  bool hres = p.mode == 3 && p.n_tiles == 1 && env_int("PL_IGEMM_NO_HRES", 0) == 0 &&
              2 * (int)p.a_halo_bytes + p.num_kb * (int)p.b_stage_bytes <= budget;
  // TMA-fed GEMM / im2col modes: when the grid is a multiple of n_tiles every CTA keeps seeing the
  // same N tile, so that filter slice can stay in smem for the kernel's lifetime (the per-tile
  // re-read of B from L2 is what bounds the short-K layers: B is 2x the A bytes at BN = 256)
  bool bres01 = (p.mode == 0 || p.mode == 1) && epi != EPI_GENERIC && env_int("PL_IGEMM_NO_BRES01", 0) == 0 &&
                tiles >= 2 * grid && p.n_tiles <= 16 && !p.mpair && !p.cta2;
  if (bres01) {
    const int g = grid - grid % p.n_tiles;
    const int bbytes = p.num_kb * (int)p.b_stage_bytes;
    if (g * 20 < grid * 19 || bbytes > env_int("PL_IGEMM_BRES_MAX_KB", 64) * 1024) bres01 = false;
    else grid = g;
  }
  if (p.tma_out) {
    int rc;
    if (p.mode == 3)  // dense band of hk output rows x Q pixels
      rc = encode_4d(&ty, p.y, p.Nout, p.Q, p.P, p.M / p.PQ, p.ldy, p.PW, p.Q, p.hk);
    else
      rc = encode_2d(&ty, p.y, (uint64_t)p.Nout, (uint64_t)p.M, (uint64_t)p.ldy, p.PW, kBM);
    if (rc) return rc;
    if (p.tma_res) {
      if (p.mode == 3)
        rc = encode_4d(&tr, e.residual, p.Nout, p.Q, p.P, p.M / p.PQ, e.ld_res, p.PW, p.Q, p.hk);
      else
        rc = encode_2d(&tr, e.residual, (uint64_t)p.Nout, (uint64_t)p.M, (uint64_t)e.ld_res, p.PW, kBM);
      if (rc) return rc;
    }
    // shared-memory policy: at least min_stages mainloop stages, then a second staging
    // buffer (the TMA store of tile i drains while tile i+1 is written), then a deeper
    // residual ring, the rest goes back to the mainloop
    const int min_stages = env_int("PL_IGEMM_MIN_STAGES", 3);
    int main_min = p.mode == 3 ? 2 * (int)p.a_halo_bytes + 4 * (int)p.b_stage_bytes : min_stages * stage_bytes;
    if (hres) main_min = 2 * (int)p.a_halo_bytes + p.num_kb * (int)p.b_stage_bytes;
    if (bres01) main_min = min_stages * (int)p.a_stage_bytes + p.num_kb * (int)p.b_stage_bytes;
    // fused residual: the output tile is written in place over the residual tile (one ring of
    // res_bufs tiles, no separate staging buffer); otherwise one or two staging buffers
    const int fixed = p.tma_res ? 2 * tile : tile;
    int left = budget - main_min - fixed;
    if (hres && left < 0) {  // the filter does not fit beside the staging buffers: stream it
      hres = false;
      main_min = 2 * (int)p.a_halo_bytes + 4 * (int)p.b_stage_bytes;
      left = budget - main_min - fixed;
    }
    if (bres01 && left < 0) {
      bres01 = false;
      grid = tiles < driver().sm_count ? tiles : driver().sm_count;  // (never a pair launch: bres01 excludes them)
      main_min = min_stages * stage_bytes;
      left = budget - main_min - fixed;
    }
    if (left < 0) left = 0;
    if (p.tma_res) {
      p.out_bufs = 0;
      int want = env_int("PL_IGEMM_RES_BUFS", 4);
      if (want > kMaxResBufs) want = kMaxResBufs;
      while (p.res_bufs < want && left >= tile) {
        ++p.res_bufs;
        left -= tile;
      }
    } else {
      p.out_bufs = 1;
      // small tiles turn over faster than a TMA store is read out: up to 4 staging buffers
      // (narrow tiles: 8 buffers, 4 of them with their stores in flight -- stem 7x7 134.7 -> 118.6 us)
      // (s2d row pairs: 16 KB tiles that turn over as fast as the 8 KB ones -- eight buffers, four stores in flight)
      int want_out = env_int("PL_IGEMM_OUT_BUFS", (tile <= 8 * 1024 || (p.s2d && p.mpair)) ? 8 : (tile <= 16 * 1024 ? 4 : 2));
      if (want_out > kMaxOutBufs) want_out = kMaxOutBufs;
      while (p.out_bufs < want_out && left >= tile) {
        ++p.out_bufs;
        left -= tile;
      }
    }
    extra = p.out_bufs * tile + (p.tma_res ? p.res_bufs * tile : 0);
    {
      const int nb = p.tma_res ? p.res_bufs : p.out_bufs;
      // deferred release pays only with many small buffers; the in-place residual ring must hand its buffers back
      // at once (its next residual load waits for them: 87 -> 182 us on r2 1x1 64->256+res with 4 in flight)
      int d = env_int("PL_STORE_DEFER", -1);
      if (d < 0) d = (!p.tma_res && nb >= 8) ? 4 : 0;
      if (d > nb - 1) d = nb - 1;
      if (d > 4) d = 4;
      p.store_defer = d < 0 ? 0 : d;
    }
  }
  p.epi_groups = 1;
  if (p.tma_out && p.sb_all) {
    const int cq = (p.mode == 2 ? kThinEpiWarps : kEpiWarps) / 4;  // warps per TMEM lane quarter
    const int per_warp_max = 16 / cq;                               // unrolled unit slots of one warp
    const int units = p.BN / 16;
    int cqn_min = (units + per_warp_max - 1) / per_warp_max;
    if (cqn_min < 1) cqn_min = 1;
    int g = cq / cqn_min;
    if (units > env_int("PL_EPI_GROUPS_MAX_UNITS", 4)) g = 1;  // measured: BN >= 128 tiles lose a few % with groups
    const int bufs = p.tma_res ? p.res_bufs : p.out_bufs;
    while (g > 1 && (g > bufs || g > p.acc_stages || (cq % g) != 0)) g >>= 1;
    // the kernel instantiates 1 group or `cq` groups.  (Two half-width groups on alternate 256-wide tiles, and on
    // 128-wide tiles with four accumulator stages, were built and measured for the epilogue-bound 1x1 expand layers:
    // 5-20 % slower, profiles/r02_epilogue.md -- both accumulator stages are then held by epilogues and the MMA of
    // the next tile cannot run ahead)
    if (g != cq || env_int("PL_EPI_GROUPS", 4) <= 1 || p.mpair) g = 1;  // (pairs: all warps on the one stage)
    p.epi_groups = g;
  }
  if (p.s2d && !hres) return PL_ERR_UNSUPPORTED;
  if ((p.n_peers > 0 || p.mc_out) && !p.tma_out) return PL_ERR_UNSUPPORTED;
  if (!p.tma_out) bres01 = false;
  {
    // measured (tools/gemm_sweep.py): bands of 16 rows for 8192^3 / 16384^3 (2621 -> 2797, 2531 -> 2880 TOP/s),
    // 8 when there are fewer than 32 tile rows; no effect at 4096^3
    int gm = env_int("PL_GROUP_M", 0);
    if (gm == 0) gm = p.m_tiles >= 32 ? 16 : 8;
    // only when B (N x K bytes) cannot stay in L2 next to a wave's A rows: with B <= 32 MB the row-major
    // order already re-uses all of it and banding loses 7 % (8192 x 4096 x 8192, 4096^3)
    const long long b_total = (long long)p.Nout * p.num_kb * p.BK;
    p.group_m = (p.mode == 0 && !bres01 && !p.s2d && gm > 1 && p.n_tiles >= 8 && p.m_tiles >= 2 * gm &&
                 (b_total >= (48ll << 20) || env_int("PL_GROUP_M", 0) > 0))
                    ? gm
                    : 1;
  }
  p.b_resident = ((p.mode == 2 && p.n_tiles == 1) || hres || bres01) ? 1 : 0;
  if (p.mode == 2 && p.b_resident && env_int("PL_THIN_NO_GROUP", 0) == 0) {
    p.thin_group = 1;
    p.a_stage_bytes = (uint32_t)p.num_kb * (uint32_t)(kBM * p.BK);  // a stage = every k-block of a tile
  }
  if (bres01) p.tx_bytes = p.a_stage_bytes;  // the stage ring carries only A blocks
  p.dbg_flags = env_int("PL_DBG_FLAGS", 0);
  p.dbg_ashift = env_int("PL_DBG_ASHIFT", 0);
  p.dbg_baseoff = env_int("PL_DBG_BASEOFF", 0);
  int stages;
  size_t ring_bytes;
  if (p.mode == 3 && hres) {
    // resident filter; what is left goes to halo buffers (deeper input prefetch)
    const int bbytes = p.num_kb * (int)p.b_stage_bytes;
    int hb = (budget - extra - bbytes) / (int)p.a_halo_bytes;
    p.halo_bufs = hb > kMaxHaloBufs ? kMaxHaloBufs : hb;
    p.a_region_bytes = (uint32_t)p.halo_bufs * p.a_halo_bytes;
    stages = 2;  // the stage ring and its barriers are unused
    ring_bytes = (size_t)p.a_region_bytes + (size_t)bbytes;
  } else if (p.mode == 3) {
    p.halo_bufs = 2;
    p.a_region_bytes = 2u * p.a_halo_bytes;
    stages = (budget - extra - (int)p.a_region_bytes) / (int)p.b_stage_bytes;
    if (stages > kMaxTmaStages) stages = kMaxTmaStages;
    ring_bytes = (size_t)p.a_region_bytes + (size_t)stages * p.b_stage_bytes;
  } else if (p.b_resident) {
    stages = (budget - extra - p.num_kb * (int)p.b_stage_bytes) / (int)p.a_stage_bytes;
    if (stages > (p.mode == 2 ? kMaxStages : kMaxTmaStages)) stages = p.mode == 2 ? kMaxStages : kMaxTmaStages;
    ring_bytes = (size_t)stages * p.a_stage_bytes + (size_t)p.num_kb * p.b_stage_bytes;
  } else {
    stages = (budget - extra) / stage_bytes;
    if (stages > (p.mode == 2 ? kMaxStages : kMaxTmaStages)) stages = p.mode == 2 ? kMaxStages : kMaxTmaStages;
    ring_bytes = (size_t)stages * stage_bytes;
  }
  if (stages < 2 || (p.mode == 2 && !p.thin_group && stages < p.num_kb)) return PL_ERR_UNSUPPORTED;
  // (Moving the MMA thread's accumulator-stage wait into the producer -- it would hold a tile's first load back
  // instead -- was measured: thin-tile layers +8...+23 %, profiles/r02_mma_thread.md: the producer has no slack either)
  p.kgroup = 1;
  if ((p.mode == 0 || p.mode == 1) && p.num_kb >= 2) {
    // group ring slots until a group's MMAs run >= ~900 cycles (the MMA thread's wait + fence + commit per group
    // costs ~450), keeping at least two groups in the ring
    const int cyc = p.BN >= 256 ? 136 : (p.BN >= 128 ? 67 : (p.BN >= 64 ? 50 : 42));
    const int per_kb = (p.BK / 32) * cyc * (p.mpair ? 2 : 1);
    // measured (profiles/r02_exp7_kgroup.txt): grouping halves the look-ahead of the ring and LOSES 5-15 % on the
    // streamed layers and GEMMs (8192^3 68 -> 58 %), small layers do not move -- so one k-block per barrier stays
    // the default; PL_KGROUP=0 asks for the automatic grouping, 2 / 4 force a group size
    int g = env_int("PL_KGROUP", 1);
    if (g <= 0) {
      g = 1;
      while (g < 4 && per_kb * g < env_int("PL_KGROUP_CYCLES", 900) && 2 * g <= p.num_kb) g *= 2;
    }
    while (g > 1 && stages / g < 2) g >>= 1;
    p.kgroup = g;
    stages -= stages % g;
  }
  p.stages = stages;
  if (p.mode != 3) p.a_region_bytes = (uint32_t)stages * p.a_stage_bytes;
  p.out_stage_off = (uint32_t)ring_bytes;
  p.res_stage_off = p.out_stage_off + (uint32_t)(p.out_bufs * tile);
  p.sb_off = (uint32_t)(ring_bytes + extra);
  const size_t smem = ring_bytes + extra + sb_bytes + 1024;
  if (p.mode == 2) {
    // thin-input gather mode: (filter rows, window words) instantiations
    const int key = p.thin_r * 16 + p.thin_sw;
    const bool relu = (epi == EPI_I8_RELU);
    if (!relu) {  // everything else (fp32 / signed / legacy) takes the generic epilogue
      p.tma_out = 0;
      p.tma_res = 0;
    }
#define PL_THIN_CASE(R_, SW_)                                                                                   \
  case R_ * 16 + SW_:                                                                                            \
    return relu ? launch_one<EPI_I8_RELU, R_, SW_>(ta, tb, ty, tr, p, grid, smem, stream)                        \
                : launch_one<EPI_GENERIC, R_, SW_>(ta, tb, ty, tr, p, grid, smem, stream);
    switch (key) {
      PL_THIN_CASE(7, 8)
      PL_THIN_CASE(5, 8)
      PL_THIN_CASE(3, 4)
      PL_THIN_CASE(3, 8)
      default: return PL_ERR_UNSUPPORTED;
    }
#undef PL_THIN_CASE
  }
  {
    const int plan[10] = {p.mode, p.BN, p.BK, p.stages | (p.kgroup << 8), p.cta2, p.mpair, p.group_m, grid, p.b_resident, epi};
    for (int i = 0; i < 10; ++i) g_last_plan[i] = plan[i];
  }
  if (p.cta2) {
    if (p.group_m > 1) {
      switch (epi) {
        case EPI_I8_RELU: return launch_one<EPI_I8_RELU, 0, 0, true, true>(ta, tb, ty, tr, p, grid, smem, stream);
        case EPI_I8_SIGNED: return launch_one<EPI_I8_SIGNED, 0, 0, true, true>(ta, tb, ty, tr, p, grid, smem, stream);
        case EPI_GENERIC: return launch_one<EPI_GENERIC, 0, 0, true, true>(ta, tb, ty, tr, p, grid, smem, stream);
        default: p.group_m = 1; break;
      }
    }
    switch (epi) {
      case EPI_I8_RELU: return launch_one<EPI_I8_RELU, 0, 0, false, true>(ta, tb, ty, tr, p, grid, smem, stream);
      case EPI_I8_RELU_RES: return launch_one<EPI_I8_RELU_RES, 0, 0, false, true>(ta, tb, ty, tr, p, grid, smem, stream);
      case EPI_I8_SIGNED: return launch_one<EPI_I8_SIGNED, 0, 0, false, true>(ta, tb, ty, tr, p, grid, smem, stream);
      default: return launch_one<EPI_GENERIC, 0, 0, false, true>(ta, tb, ty, tr, p, grid, smem, stream);
    }
  }
  if (p.group_m > 1) {  // large GEMMs only (mode 0, streamed B): their own instantiations
    switch (epi) {
      case EPI_I8_RELU: return launch_one<EPI_I8_RELU, 0, 0, true>(ta, tb, ty, tr, p, grid, smem, stream);
      case EPI_I8_SIGNED: return launch_one<EPI_I8_SIGNED, 0, 0, true>(ta, tb, ty, tr, p, grid, smem, stream);
      case EPI_GENERIC: return launch_one<EPI_GENERIC, 0, 0, true>(ta, tb, ty, tr, p, grid, smem, stream);
      default: p.group_m = 1; break;  // (fused residual: row-major order)
    }
  }
  switch (epi) {
    case EPI_I8_RELU: return launch_one<EPI_I8_RELU>(ta, tb, ty, tr, p, grid, smem, stream);
    case EPI_I8_RELU_RES: return launch_one<EPI_I8_RELU_RES>(ta, tb, ty, tr, p, grid, smem, stream);
    case EPI_I8_SIGNED: return launch_one<EPI_I8_SIGNED>(ta, tb, ty, tr, p, grid, smem, stream);
    default: return launch_one<EPI_GENERIC>(ta, tb, ty, tr, p, grid, smem, stream);
  }
}

thread_local int g_last_plan[10] = {};

int max_pair_clusters() {
  static int cached[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  int& c = cached[dev & 63];
  if (c == 0) {
    auto kern = igemm_i8_kernel<EPI_I8_RELU, 0, 0, false, true>;
    c = -1;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(221 * 1024)) == cudaSuccess) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(2 * driver().sm_count);
      cfg.blockDim = dim3(kThreads);
      cfg.dynamicSmemBytes = 221 * 1024;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2;
      at[0].val.clusterDim.y = 1;
      at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0) c = n;
    }
    cudaGetLastError();
  }
  return c > 0 ? c : 0;
}

int check_epi(const pl_epilogue* e) {
  if (!e) return PL_ERR_ARG;
  if (e->out_dtype != PL_OUT_INT8 && e->out_dtype != PL_OUT_FP32 && e->out_dtype != PL_OUT_INT32)
    return PL_ERR_ARG;
  if (e->out_dtype != PL_OUT_INT32 && !e->scale) return PL_ERR_ARG;
  return PL_OK;
}

// p.BK holds the packing K-block of the reduction dim (pl_conv_block_k); the launch may
// use a smaller one (the packed rows are plain zero-padded K-major bytes)
int common_setup(IgemmParams& p, long long M, int Nout, void* y, int ldy, const pl_epilogue* e, int Kc,
                 int taps) {
  if (M <= 0 || Nout <= 0 || M > 0x7fffffffLL) return PL_ERR_ARG;
  const int cpad = (Kc + p.BK - 1) / p.BK * p.BK;
  p.M = (int)M;
  p.Nout = Nout;
  p.m_tiles = p.halo_m_tiles > 0 ? p.halo_m_tiles : (int)((M + kBM - 1) / kBM);
  p.BN = pick_bn(Nout, p.m_tiles, driver().sm_count);
  if (p.s2d) {  // the filter must be resident: one N tile, never narrowed to fill the grid
    p.BN = 16;
    while (p.BN < Nout && p.BN < 256) p.BN <<= 1;
  }
  if (p.halo_m_tiles > 0 && !p.s2d && Nout <= 256 && env_int("PL_HALO_KEEP_NTILE", 1)) {
    // halo launches that do not fill the SMs: narrowing the tile multiplies the CTAs, but every one of them still
    // loads the whole halo tile, and the filter stops being resident -- the MMA thread then walks R*S filter
    // stages (wait + commit, ~450 cycles each: 4000 of the 6500 cycles of a one-tile 3x3 launch, CTA-0 timeline
    // in profiles/r02_latency.md) instead of issuing all taps in one run.  Keep ONE N tile while its packed
    // filter fits beside two halo tiles
    int bn1 = 16;
    while (bn1 < Nout) bn1 <<= 1;
    const long long filt = (long long)taps * cpad * bn1;
    if (bn1 > p.BN && filt <= 150 * 1024) p.BN = bn1;
  }
  {
    // a 256-wide tile whose filter slice (K_red x 256 bytes) cannot stay resident in shared memory, while a 128-wide
    // one (<= 64 KB) can: the resident slice wins (r5 1x1 512 -> 2048 + residual: 25.8 -> 21.6 us at b256)
    const long long kred = (long long)taps * cpad;
    const int nt128 = (Nout + 127) / 128;
    if (p.BN == 256 && p.halo_m_tiles == 0 && !p.s2d && p.mode != 2 && e->out_dtype == PL_OUT_INT8 && kred * 256 > 64 * 1024 && kred * 128 <= 64 * 1024 &&
        nt128 <= 16 && (long long)p.m_tiles * nt128 >= 2ll * driver().sm_count && env_int("PL_NO_BN128_RES", 0) == 0)
      p.BN = 128;
  }
  if (e->residual) {
    // fused-residual launches are bandwidth bound and need smem for the residual ring and
    // the staging buffers: thin reductions take a finer K-block so that three mainloop
    // stages still fit beside them (measured on B200: tools/gpu_sweep_res.sh)
    const int cap = env_int("PL_IGEMM_RES_BN", 256);
    while (p.BN > cap && p.BN > 16) p.BN >>= 1;
    const int bk_cap = env_int("PL_IGEMM_RES_BK", cpad <= 128 ? 64 : 128);
    while (p.BK > bk_cap && p.BK > 32) p.BK >>= 1;
  }
  p.cblocks = cpad / p.BK;
  p.num_kb = taps * p.cblocks;
  p.n_tiles = (Nout + p.BN - 1) / p.BN;
  p.a_stage_bytes = (uint32_t)(kBM * p.BK);
  p.b_stage_bytes = (uint32_t)((p.BN * p.BK + 1023) / 1024 * 1024);
  p.tx_bytes = (uint32_t)(kBM * p.BK + p.BN * p.BK);
  p.acc_stride = (uint32_t)(p.BN < 32 ? 32 : p.BN);
  p.acc_stages = 512 / (int)p.acc_stride;
  if (p.acc_stages > kMaxAccStages) p.acc_stages = kMaxAccStages;
  p.tmem_cols = pow2_cols(p.acc_stages * (int)p.acc_stride);
  p.y = y;
  p.ldy = ldy;
  p.epi = make_epi(e, Nout);
  const int eb = e->out_dtype == PL_OUT_INT8 ? 1 : 4;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(y) & 15) == 0) && (((long long)ldy * eb) % 16 == 0);
  p.res_vec_ok = e->residual && ((reinterpret_cast<uintptr_t>(e->residual) & 15) == 0) && (e->ld_res % 16 == 0);
  return PL_OK;
}

}  // namespace
}  // namespace pl

// ============================================================================
// C-ABI
// ============================================================================
using namespace pl;
namespace pl {
int launch_gemv_i8(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, int ldc, int M, int N, int K,
                   const pl_epilogue* e, pl_stream_t stream);
}

extern "C" int pl_conv_block_k(int C) { return C > 64 ? 128 : (C > 32 ? 64 : 32); }

extern "C" size_t pl_conv_packed_weight_bytes(int K, int R, int S, int C) {
  const int bk = pl_conv_block_k(C);
  const int cpad = (C + bk - 1) / bk * bk;
  return (size_t)K * R * S * cpad;
}

extern "C" int pl_pack_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_packed, int K, int R, int S,
                                      int C, pl_stream_t stream) {
  if (!w_ohwi || !w_packed || K <= 0 || R <= 0 || S <= 0 || C <= 0) return PL_ERR_ARG;
  const int bk = pl_conv_block_k(C);
  const int cpad = (C + bk - 1) / bk * bk;
  const long long total = (long long)K * R * S * cpad;
  const int blocks = (int)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
  pack_conv_weight_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_ohwi, w_packed, K, R * S, C, cpad);
  ++g_launches;
  return launch_status();
}

// one dense (groups == 1) launch; allow_halo = false keeps stride-1 windows on the im2col TMA path
static int conv2d_dense(const int8_t* x, const int8_t* w_packed, void* y, const pl_conv_desc* d, const pl_epilogue* e,
                        pl_stream_t stream, bool allow_halo, bool* used_halo) {
  *used_halo = false;
  int rc;
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->C <= 0 || d->K <= 0 || d->R <= 0 || d->S <= 0 ||
      d->stride_h <= 0 || d->stride_w <= 0 || d->dil_h <= 0 || d->dil_w <= 0 || d->ldx < d->C ||
      d->ldy < d->K)
    return PL_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(w_packed) & 15) ||
      (d->ldx % 16))
    return PL_ERR_ALIGN;
  const int P = pl_conv_out_size(d->H, d->R, d->dil_h, d->pad_t, d->pad_b, d->stride_h);
  const int Q = pl_conv_out_size(d->W, d->S, d->dil_w, d->pad_l, d->pad_r, d->stride_w);
  if (P <= 0 || Q <= 0) return PL_ERR_ARG;
  const int up_w = d->pad_r - (d->S - 1) * d->dil_w, up_h = d->pad_b - (d->R - 1) * d->dil_h;
  if (d->pad_l > 128 || d->pad_t > 128 || up_w < -128 || up_w > 127 || up_h < -128 || up_h > 127 ||
      (d->S - 1) * d->dil_w > 255 || (d->R - 1) * d->dil_h > 255 || d->stride_w > 8 || d->stride_h > 8)
    return PL_ERR_UNSUPPORTED;

  IgemmParams p{};
  p.BK = pl_conv_block_k(d->C);
  const int cpad = (d->C + p.BK - 1) / p.BK * p.BK;
  const bool as_gemm = d->R == 1 && d->S == 1 && d->stride_h == 1 && d->stride_w == 1 &&
                       d->pad_t == 0 && d->pad_b == 0 && d->pad_l == 0 && d->pad_r == 0;
  // halo reuse (mode 3): stride-1, undilated windows whose padded-grid bands fill the 128-row
  // tile well enough; everything else keeps the im2col TMA path
  const int WP = Q + d->S - 1;
  const int hk = WP <= kBM ? kBM / WP : 0;
  bool halo = allow_halo && !as_gemm && d->stride_h == 1 && d->stride_w == 1 && d->dil_h == 1 && d->dil_w == 1 && hk >= 1 &&
              WP <= 256 && hk + d->R - 1 <= 256 && Q <= 256 && env_int("PL_IGEMM_NO_HALO", 0) == 0 &&
              d->W + d->pad_l + d->pad_r == WP;
  int bands = 0;
  if (halo) {
    bands = (P + hk - 1) / hk;
    const double eff = (double)P * Q / ((double)bands * kBM);  // useful rows per MMA tile
    if (eff < 0.6) halo = false;
  }
  if (halo) p.halo_m_tiles = d->N * bands;
  *used_halo = halo;
  rc = common_setup(p, (long long)d->N * P * Q, d->K, y, d->ldy, e, d->C, d->R * d->S);
  if (rc) return rc;
  CUtensorMap ta, tb;
  if (halo) {
    p.mode = 3;
    p.Q = Q;
    p.P = P;
    p.PQ = P * Q;
    p.S = d->S;
    p.R = d->R;
    p.pt = d->pad_t;
    p.pl_ = d->pad_l;
    p.hk = hk;
    p.WP = WP;
    p.bands = bands;
    p.dBands = make_fastdiv((uint32_t)bands);
    p.dWP = make_fastdiv((uint32_t)WP);
    const int rows_loaded = (hk + d->R - 1) * WP;
    const int rows_read = kBM + (d->R - 1) * WP + d->S - 1;  // furthest row an MMA descriptor touches
    const int rows_alloc = rows_loaded > rows_read ? rows_loaded : rows_read;
    p.a_halo_tx = (uint32_t)(rows_loaded * p.BK);
    p.a_halo_bytes = (uint32_t)((rows_alloc * p.BK + 1023) / 1024 * 1024);
    p.tx_bytes = (uint32_t)(p.BN * p.BK);  // the stage ring carries only filter blocks
    rc = encode_4d(&ta, x, d->C, d->W, d->H, d->N, d->ldx, p.BK, WP, hk + d->R - 1);
  } else if (as_gemm) {
    p.mode = 0;
    rc = encode_2d(&ta, x, (uint64_t)d->C, (uint64_t)p.M, (uint64_t)d->ldx, p.BK, kBM);
  } else {
    p.mode = 1;
    p.Q = Q;
    p.PQ = P * Q;
    p.dQ = make_fastdiv((uint32_t)Q);
    p.dPQ = make_fastdiv((uint32_t)(P * Q));
    p.S = d->S;
    p.sh = d->stride_h;
    p.sw = d->stride_w;
    p.pt = d->pad_t;
    p.pl_ = d->pad_l;
    p.dh = d->dil_h;
    p.dw = d->dil_w;
    rc = encode_im2col(&ta, x, d, p.BK);
  }
  if (rc) return rc;
  const uint64_t wrow = (uint64_t)d->R * d->S * cpad;
  rc = encode_2d(&tb, w_packed, wrow, (uint64_t)d->K, wrow, p.BK, p.BN);
  if (rc) return rc;
  const BSpec bs{w_packed, wrow, (uint64_t)d->K, wrow};
  return launch_igemm(ta, tb, p, stream, &bs);
}

extern "C" int pl_conv2d_i8_nhwc(const int8_t* x, const int8_t* w_packed, void* y,
                                 const pl_conv_desc* d, const pl_epilogue* e, pl_stream_t stream) {
  if (!x || !w_packed || !y || !d) return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if (d->groups < 1 || d->C % d->groups || d->K % d->groups) return PL_ERR_ARG;
  auto one = [&](const int8_t* xg, const int8_t* wg, void* yg, const pl_conv_desc* dg, const pl_epilogue* eg) {
    bool used_halo = false;
    int r = conv2d_dense(xg, wg, yg, dg, eg, stream, true, &used_halo);
    // halo mode is chosen from the geometry; when its two input buffers + the filter ring do not fit the
    // shared memory (e.g. 5x5 / 7x7 stride 1 with C >= 128 on wide images) the im2col path takes the layer
    if (r == PL_ERR_UNSUPPORTED && used_halo) r = conv2d_dense(xg, wg, yg, dg, eg, stream, false, &used_halo);
    return r;
  };
  if (d->groups == 1) return one(x, w_packed, y, d, e);
  // 1 < groups < C (lite/kernels/arm/conv_gemmlike.cc:84-264 loops over the groups the same way): group g reads
  // the channel slice [g * C/G, +C/G) of every pixel (pointer offset + the full pitch), the filter rows
  // [g * K/G, +K/G) of the packed filter (packed with C = C/G) and writes the output slice [g * K/G, +K/G)
  const int G = d->groups, Cg = d->C / G, Kg = d->K / G;
  if (Cg % 16) return PL_ERR_UNSUPPORTED;  // every group's slice must start on a 16-byte boundary (TMA)
  if (e->residual) return PL_ERR_UNSUPPORTED;
  const int bk = pl_conv_block_k(Cg);
  const size_t wrow = (size_t)d->R * d->S * ((Cg + bk - 1) / bk * bk);
  const size_t yb = e->out_dtype == PL_OUT_INT8 ? 1 : 4;
  for (int g = 0; g < G; ++g) {
    pl_conv_desc dg = *d;
    dg.groups = 1;
    dg.C = Cg;
    dg.K = Kg;
    pl_epilogue eg = *e;
    if (eg.scale) eg.scale += (size_t)g * Kg;
    if (eg.bias) eg.bias += (size_t)g * Kg;
    rc = one(x + (size_t)g * Cg, w_packed + (size_t)g * Kg * wrow,
             reinterpret_cast<char*>(y) + (size_t)g * Kg * yb, &dg, &eg);
    if (rc) return rc;
  }
  return PL_OK;
}

// ---- thin first layers: direct conv on one-word-per-pixel input ------------------
static int thin_sw(int S) { return S <= 4 ? 4 : 8; }
static int thin_shift(int S) { return S < thin_sw(S) ? 1 : 0; }  // spare window word in front (see IgemmParams)
static int thin_nkb(int R, int S) {
  const int rpk = 8 / thin_sw(S);
  return (R + rpk - 1) / rpk;
}

extern "C" int pl_thin_conv_supported(int R, int S, int C, int dil_h, int dil_w) {
  if (C < 1 || C > 4 || dil_h != 1 || dil_w != 1 || S < 1 || S > 8) return 0;
  const int sw = thin_sw(S);
  return (R == 7 && sw == 8) || (R == 5 && sw == 8) || (R == 3 && sw == 4) || (R == 3 && sw == 8);
}

extern "C" size_t pl_thin_conv_packed_weight_bytes(int K, int R, int S) {
  if (K <= 0 || R <= 0 || S <= 0 || S > 8) return 0;
  return (size_t)K * thin_nkb(R, S) * 32;
}

extern "C" int pl_pack_thin_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_thin, int K, int R, int S, int C,
                                           pl_stream_t stream) {
  if (!w_ohwi || !w_thin || K <= 0 || !pl_thin_conv_supported(R, S, C, 1, 1)) return PL_ERR_ARG;
  const int row_bytes = thin_nkb(R, S) * 32;
  const long long total = (long long)K * row_bytes;
  const int blocks = (int)((total + 255) / 256 > 1024 ? 1024 : (total + 255) / 256);
  pack_thin_weight_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_ohwi, w_thin, K, R, S, C, thin_sw(S),
                                                                    row_bytes, thin_shift(S));
  ++g_launches;
  return launch_status();
}

extern "C" int pl_thin_conv2d_i8_nhwc(const int8_t* x, const int8_t* w_thin, void* y, const pl_conv_desc* d,
                                      const pl_epilogue* e, pl_stream_t stream) {
  if (!x || !w_thin || !y || !d) return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if (d->groups != 1 || e->residual) return PL_ERR_UNSUPPORTED;
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->K <= 0 || d->stride_h <= 0 || d->stride_w <= 0 || d->ldy < d->K)
    return PL_ERR_ARG;
  if (!pl_thin_conv_supported(d->R, d->S, d->C, d->dil_h, d->dil_w)) return PL_ERR_UNSUPPORTED;
  if (d->ldx != 4 || (reinterpret_cast<uintptr_t>(x) & 3) || (reinterpret_cast<uintptr_t>(w_thin) & 15))
    return PL_ERR_ALIGN;
  const int P = pl_conv_out_size(d->H, d->R, 1, d->pad_t, d->pad_b, d->stride_h);
  const int Q = pl_conv_out_size(d->W, d->S, 1, d->pad_l, d->pad_r, d->stride_w);
  if (P <= 0 || Q <= 0) return PL_ERR_ARG;
  if ((long long)d->N * P * Q >= (1ll << 31) || (long long)d->H * d->W >= (1ll << 31)) return PL_ERR_UNSUPPORTED;
  IgemmParams p{};
  p.mode = 2;
  p.BK = 32;
  const int nkb = thin_nkb(d->R, d->S);
  rc = common_setup(p, (long long)d->N * P * Q, d->K, y, d->ldy, e, nkb * 32, 1);
  if (rc) return rc;
  p.xw = reinterpret_cast<const uint32_t*>(x);
  p.H = d->H;
  p.W = d->W;
  p.Q = Q;
  p.PQ = P * Q;
  p.sh = d->stride_h;
  p.sw = d->stride_w;
  p.pt = d->pad_t;
  p.pl_ = d->pad_l;
  p.dQ = make_fastdiv((uint32_t)Q);
  p.dPQ = make_fastdiv((uint32_t)(P * Q));
  p.thin_r = d->R;
  p.thin_sw = thin_sw(d->S);
  p.thin_shift = thin_shift(d->S);
  p.thin_pairs = (d->stride_w % 2 == 0) && ((d->pad_l + p.thin_shift) % 2 == 0) && (d->W % 2 == 0) &&
                 ((reinterpret_cast<uintptr_t>(x) & 7) == 0) && env_int("PL_THIN_NO_PAIRS", 0) == 0;
  p.tx_bytes = (uint32_t)(p.BN * p.BK);  // only the filter tile arrives by TMA
  CUtensorMap tb;
  rc = encode_2d(&tb, w_thin, (uint64_t)nkb * 32, (uint64_t)d->K, (uint64_t)nkb * 32, p.BK, p.BN);
  if (rc) return rc;
  return launch_igemm(tb, tb, p, stream);
}

// ---- stride-2 thin stems in space-to-depth form -------------------------------------------------
extern "C" pl_s2d_geometry pl_s2d_conv_geometry(int N, int H, int W, int C, int R, int S, int stride_h,
                                                int stride_w, int pad_t, int pad_b, int pad_l, int pad_r,
                                                int dil_h, int dil_w) {
  pl_s2d_geometry g{};
  if (N <= 0 || H <= 0 || W <= 0 || C < 1 || C > 4 || R != S || R < 1 || stride_h != 2 || stride_w != 2 ||
      dil_h != 1 || dil_w != 1 || (H & 1) || (W & 1) || pad_t != pad_l || pad_t < 0 || pad_b < 0 || pad_r < 0)
    return g;
  g.P = pl_conv_out_size(H, R, 1, pad_t, pad_b, 2);
  g.Q = pl_conv_out_size(W, S, 1, pad_l, pad_r, 2);
  if (g.P <= 0 || g.Q <= 0 || g.Q > kBM) return g;
  g.pad_s = (pad_t + 1) / 2;
  g.shift = 2 * g.pad_s - pad_t;
  g.TA = (g.shift + R + 1) / 2;
  g.TB = (g.TA + 1) / 2 * 2;
  if (g.TA > 4) return g;
  const int Hs = H / 2, Ws = W / 2;
  // last s2d row / pixel any window touches
  const int last_r = (g.P - 1) + g.TA - 1, last_c = (g.Q - 1) + g.TB - 1;  // in padded coordinates
  g.HP = last_r + 1 > Hs + g.pad_s ? last_r + 1 : Hs + g.pad_s;
  g.WS = last_c + 1 > Ws + g.pad_s ? last_c + 1 : Ws + g.pad_s;
  // a tile always fetches 128 + TB - 1 pixels from the start of a row: slack after the last row
  g.bytes = ((size_t)N * g.HP * g.WS + (size_t)(kBM + g.TB)) * 16;
  g.supported = 1;
  return g;
}

extern "C" size_t pl_s2d_conv_packed_weight_bytes(int K, const pl_s2d_geometry* g) {
  if (K <= 0 || !g || !g->supported) return 0;
  return (size_t)K * g->TA * g->TB * 16;
}

extern "C" int pl_pack_s2d_conv_weight_i8(const int8_t* w_ohwi, int8_t* w_s2d, int K, int R, int S, int C,
                                          const pl_s2d_geometry* g, pl_stream_t stream) {
  if (!w_ohwi || !w_s2d || K <= 0 || !g || !g->supported || C < 1 || C > 4) return PL_ERR_ARG;
  const long long total = (long long)K * g->TA * g->TB * 16;
  const int blocks = (int)((total + 255) / 256 > 1024 ? 1024 : (total + 255) / 256);
  pack_s2d_weight_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_ohwi, w_s2d, K, R, S, C, g->TA, g->TB, g->shift);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_s2d_conv2d_i8(const int8_t* x_s2d, const int8_t* w_s2d, void* y, const pl_conv_desc* d,
                                const pl_epilogue* e, pl_stream_t stream) {
  if (!x_s2d || !w_s2d || !y || !d) return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if (d->groups != 1 || e->residual) return PL_ERR_UNSUPPORTED;
  if (d->K <= 0 || d->ldy < d->K) return PL_ERR_ARG;
  const pl_s2d_geometry g = pl_s2d_conv_geometry(d->N, d->H, d->W, d->C, d->R, d->S, d->stride_h, d->stride_w,
                                                 d->pad_t, d->pad_b, d->pad_l, d->pad_r, d->dil_h, d->dil_w);
  if (!g.supported) return PL_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(x_s2d) & 15) || (reinterpret_cast<uintptr_t>(w_s2d) & 15)) return PL_ERR_ALIGN;
  if ((long long)d->N * g.P * g.Q >= (1ll << 31)) return PL_ERR_UNSUPPORTED;
  IgemmParams p{};
  p.mode = 3;
  p.s2d = 1;
  p.BK = 32;
  const int nkb = g.TA * g.TB / 2;
  p.halo_m_tiles = d->N * g.P;  // one output row per tile
  rc = common_setup(p, (long long)d->N * g.P * g.Q, d->K, y, d->ldy, e, nkb * 32, 1);
  if (rc) return rc;
  if (p.n_tiles != 1) return PL_ERR_UNSUPPORTED;  // the filter must be resident (K <= 256 outputs)
  p.Q = g.Q;
  p.P = g.P;
  p.PQ = g.P * g.Q;
  p.S = 1;
  p.R = 1;
  p.hk = 1;
  p.WP = kBM;
  p.bands = g.P;
  p.dBands = make_fastdiv((uint32_t)g.P);
  p.dWP = make_fastdiv((uint32_t)kBM);
  p.s2d_ta = g.TA;
  p.s2d_kpr = g.TB / 2;
  p.s2d_span_bytes = (uint32_t)((kBM + g.TB - 1) * 16);
  p.s2d_span_pitch = (p.s2d_span_bytes + 127u) / 128u * 128u;
  p.s2d_row_bytes = (long long)g.WS * 16;
  p.s2d_img_bytes = (long long)g.HP * g.WS * 16;
  p.xs2d = x_s2d;
  p.a_halo_tx = (uint32_t)g.TA * p.s2d_span_bytes;
  p.a_halo_bytes = ((uint32_t)g.TA * p.s2d_span_pitch + 1023u) / 1024u * 1024u;
  {
    // Row pairs: two output rows per tile (TA + 1 spans, two accumulators, one hand-off).  A stem tile is 2-8 MMAs
    // behind ~1000 cycles of per-tile work in every single-thread role (profiles/r02_mma_thread.md): a pair halves
    // that cost per row.  Needs the staged int8 epilogues (the pair leaves through one 4-D TMA store of two rows).
    const bool fast = e->out_dtype == PL_OUT_INT8 && e->round_mode == PL_ROUND_ARM && p.vec_ok && e->scale &&
                      tma_epilogue_enabled() &&
                      (e->act == PL_ACT_NONE || e->act == PL_ACT_RELU || (e->act == PL_ACT_RELU6 && e->act_arg > 0.f));
    const bool known = (g.TA == 4 && g.TB == 4) || (g.TA == 2 && g.TB == 2);  // the two unrolled MMA runs (7x7 / 3x3 stems)
    // measured at b256: the 3x3 stem (BN = 32, 2 MMAs per row) 109.3 -> 92.8 us, the 7x7 stem (BN = 64, 8 MMAs per row)
    // 113.1 -> 114.4: pairs where a row is at most 32 channels wide (PL_S2D_PAIR: 0 never, 1 that rule, 2 always)
    const int pair_knob = env_int("PL_S2D_PAIR", 1);
    if (pair_knob > 0 && (pair_knob >= 2 || p.BN <= 32) && fast && known && g.P % 2 == 0 && g.Q % 8 == 0 && p.BN <= 128) {
      p.mpair = 1;
      p.hk = 2;
      p.halo_m_tiles = d->N * (g.P / 2);
      p.m_tiles = p.halo_m_tiles;
      p.bands = g.P / 2;
      p.dBands = make_fastdiv((uint32_t)(g.P / 2));
      p.a_halo_tx = (uint32_t)(g.TA + 1) * p.s2d_span_bytes;
      p.a_halo_bytes = ((uint32_t)(g.TA + 1) * p.s2d_span_pitch + 1023u) / 1024u * 1024u;
      p.acc_stride = 2u * (uint32_t)p.BN;
      p.acc_stages = 512 / (int)p.acc_stride;
      if (p.acc_stages > kMaxAccStages) p.acc_stages = kMaxAccStages;
      p.tmem_cols = pow2_cols(p.acc_stages * (int)p.acc_stride);
      p.pair_sbase = (uint32_t)(g.Q * (p.BN < 128 ? p.BN : 128));  // Q staged rows of PW bytes
    }
  }
  p.tx_bytes = (uint32_t)(p.BN * p.BK);
  CUtensorMap tb;
  rc = encode_2d(&tb, w_s2d, (uint64_t)nkb * 32, (uint64_t)d->K, (uint64_t)nkb * 32, p.BK, p.BN);
  if (rc) return rc;
  return launch_igemm(tb, tb, p, stream);
}

// long-K GEMM: switch the launch to 256 x 256 tile pairs (IgemmParams::mpair) when it qualifies
static void maybe_enable_mpair(IgemmParams& p, const pl_epilogue* e, int K, bool peers) {
  // pairs leave through the staged int8 epilogues + TMA stores only
  const bool fast = e->out_dtype == PL_OUT_INT8 && e->round_mode == PL_ROUND_ARM && p.vec_ok && e->scale &&
                    !e->residual && tma_epilogue_enabled() &&
                    (e->act == PL_ACT_NONE || e->act == PL_ACT_RELU || (e->act == PL_ACT_RELU6 && e->act_arg > 0.f));
  // fp32 out (act none / relu): the pair stores straight from registers (IgemmParams::f32_fast), no staging
  const bool f32 = !peers && e->out_dtype == PL_OUT_FP32 && e->round_mode == PL_ROUND_ARM && p.vec_ok && e->scale &&
                   !e->residual && (e->act == PL_ACT_NONE || e->act == PL_ACT_RELU) && env_int("PL_NO_F32_FAST", 0) == 0 &&
                   env_int("PL_MPAIR_F32", 1) != 0;
  if ((!fast && !f32) || p.mode != 0 || p.BN != 256 || p.m_tiles < 2 || env_int("PL_NO_MPAIR", 0)) return;
  // CTA pairs move the same bytes per MAC as a tile pair and keep two accumulator stages (the epilogue stays
  // hidden): 8192^3 62.7 -> 70.3 % of the tensor peak, both together 57 % (BK = 64 rows halve the TMA rate)
  // (from K = 16384 the exposed epilogue of a 512 x 256 cluster tile is small enough: 16384^3 65.8 -> 70.0 %)
  const bool with_pairs = want_cta_pairs(p);
  if (with_pairs && K < env_int("PL_CTA2_MPAIR_MIN_K", 16384)) return;
  if (!env_int("PL_MPAIR_FORCE", 0)) {
    // the pair has a single accumulator stage (its epilogue is exposed), so K must be long enough
    // to hide it, and the coarser work unit must not cost a wave
    const long long sms = driver().sm_count;
    const long long tiles = (long long)p.m_tiles * p.n_tiles, pairs = (long long)((p.m_tiles + 1) / 2) * p.n_tiles;
    const long long w1 = (tiles + sms - 1) / sms, w2 = 2 * ((pairs + sms - 1) / sms);
    // (the exposed fp32 epilogue of a pair is twice as long: 4096^3 loses 12 %, 8192^3 gains 12 %)
    const int min_k = fast ? env_int("PL_MPAIR_MIN_K", 4096) : env_int("PL_MPAIR_MIN_K_F32", 8192);
    if (K < min_k || tiles < 2 * sms || w2 > w1) return;
  }
  p.mpair = 1;
  p.BK = with_pairs ? 128 : 64;  // 32 KB stages: four of them beside the 64 KB staging buffer (pairs: 48 KB stages)
  p.cblocks = (K + p.BK - 1) / p.BK;  // the K tail is zero-filled by TMA
  p.num_kb = p.cblocks;
  p.m_tiles = (p.m_tiles + 1) / 2;
  p.a_stage_bytes = (uint32_t)(2 * kBM * p.BK);
  p.b_stage_bytes = (uint32_t)(p.BN * p.BK);
  p.tx_bytes = p.a_stage_bytes + p.b_stage_bytes;
  p.acc_stride = 2u * (uint32_t)p.BN;
  p.acc_stages = 1;
  p.tmem_cols = 512;
}

// N-split GEMM whose output tiles are also stored into peer GPUs' C buffers (see IgemmParams::peer_maps)
extern "C" int pl_gemm_i8_peers(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, void* const* C_peers,
                                int n_peers, void* maps_dev, int ldc, int M, int N, int K, const pl_epilogue* e,
                                pl_stream_t stream) {
  // C_peers == NULL with n_peers > 0: the maps in maps_dev were written by pl_gemm_i8_peers_prepare
  // (the launch is then one kernel, nothing is encoded or copied: capturable in a CUDA graph)
  if (!A || !Bt || !C || M <= 0 || N <= 0 || K <= 0 || lda < K || ldb < K || ldc < N || n_peers < 0 || n_peers > 15 ||
      (n_peers > 0 && !maps_dev))
    return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(Bt) & 15) || (lda % 16) || (ldb % 16) ||
      (reinterpret_cast<uintptr_t>(maps_dev) & 63))
    return PL_ERR_ALIGN;
  IgemmParams p{};
  p.BK = pl_conv_block_k(K);
  p.mode = 0;
  rc = common_setup(p, M, N, C, ldc, e, K, 1);
  if (rc) return rc;
  maybe_enable_mpair(p, e, K, n_peers > 0);
  CUtensorMap ta, tb;
  rc = encode_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, p.BK, kBM << p.mpair);
  if (rc) return rc;
  rc = encode_2d(&tb, Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, p.BK, p.BN);
  if (rc) return rc;
  if (n_peers > 0) {
    // the peer stores exist only on the TMA-store epilogues (int8 out): same tile geometry as the local map
    if (e->out_dtype != PL_OUT_INT8 || !p.vec_ok) return PL_ERR_UNSUPPORTED;
    if (C_peers) {
      const int PW = p.BN < 128 ? p.BN : 128;
      CUtensorMap maps[15];
      for (int i = 0; i < n_peers; ++i) {
        if (!C_peers[i] || (reinterpret_cast<uintptr_t>(C_peers[i]) & 15)) return PL_ERR_ALIGN;
        rc = encode_2d(&maps[i], C_peers[i], (uint64_t)N, (uint64_t)M, (uint64_t)ldc, PW, kBM);
        if (rc) return rc;
      }
      PL_CUDA_OK(cudaMemcpyAsync(maps_dev, maps, sizeof(CUtensorMap) * n_peers, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    }
    p.peer_maps = reinterpret_cast<const CUtensorMap*>(maps_dev);
    p.n_peers = n_peers;
  }
  const BSpec bs{Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb};
  return launch_igemm(ta, tb, p, stream, &bs);
}

// N-split GEMM whose output tiles leave through an NVSwitch multicast address (IgemmParams::mc_out): C is this
// rank's own view of its column block (alignment checks, never written by this call), C_mc the multicast address
// of the same block -- every rank bound to the multicast object, this one included, receives the tiles
extern "C" int pl_gemm_i8_mcast(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, void* C_mc, int ldc,
                                int M, int N, int K, const pl_epilogue* e, pl_stream_t stream) {
  if (!A || !Bt || !C || !C_mc || M <= 0 || N <= 0 || K <= 0 || lda < K || ldb < K || ldc < N) return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(Bt) & 15) || (lda % 16) || (ldb % 16) ||
      (reinterpret_cast<uintptr_t>(C_mc) & 15) || (ldc % 16))
    return PL_ERR_ALIGN;
  // (fp32 results: multimem.st straight from the epilogue registers -- 16-byte pieces of 32 different rows per warp
  // store -- measured 10x slower than the GEMM itself; they take pl_mcast_copy_2d after a plain pl_gemm_i8 instead)
  if (e->out_dtype != PL_OUT_INT8 || e->residual) return PL_ERR_UNSUPPORTED;
  IgemmParams p{};
  p.BK = pl_conv_block_k(K);
  p.mode = 0;
  rc = common_setup(p, M, N, C, ldc, e, K, 1);
  if (rc) return rc;
  if (!p.vec_ok) return PL_ERR_UNSUPPORTED;
  maybe_enable_mpair(p, e, K, true);
  CUtensorMap ta, tb;
  rc = encode_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, p.BK, kBM << p.mpair);
  if (rc) return rc;
  rc = encode_2d(&tb, Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, p.BK, p.BN);
  if (rc) return rc;
  p.mc_out = reinterpret_cast<int8_t*>(C_mc);
  const BSpec bs{Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb};
  rc = launch_igemm(ta, tb, p, stream, &bs);
  return rc;
}

// rows x row_bytes block: local memory -> the multicast address of the same block (coalesced 16-byte multimem.st:
// 512 contiguous bytes per warp store).  The all-gather step of an N-split GEMM whose result type has no staged
// epilogue (fp32): every rank broadcasts its own column block, the switch writes it into all ranks' results.
__global__ void __launch_bounds__(256) mcast_copy_kernel(const uint8_t* __restrict__ src, long long ld_src, uint8_t* mc,
                                                        long long ld_dst, long long rows, int chunks_per_row) {
  pdl_launch_dependents();
  pdl_wait();
  const long long total = rows * chunks_per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / chunks_per_row;
    const long long c = i - r * chunks_per_row;
    const uint4 v = *reinterpret_cast<const uint4*>(src + r * ld_src + c * 16);
    multimem_st_v4(mc + r * ld_dst + c * 16, v);
  }
}

extern "C" int pl_mcast_copy_2d(const void* src, int64_t ld_src_bytes, void* dst_mc, int64_t ld_dst_bytes, int64_t rows,
                                int64_t row_bytes, pl_stream_t stream) {
  if (!src || !dst_mc || rows <= 0 || row_bytes <= 0 || ld_src_bytes < row_bytes || ld_dst_bytes < row_bytes)
    return PL_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst_mc) & 15) || (row_bytes & 15) ||
      (ld_src_bytes & 15) || (ld_dst_bytes & 15))
    return PL_ERR_ALIGN;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  const long long total = rows * (row_bytes / 16);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)driver().sm_count * 8;
  if (blocks > cap) blocks = cap;
  launch_pdl(mcast_copy_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream,
             reinterpret_cast<const uint8_t*>(src), (long long)ld_src_bytes, reinterpret_cast<uint8_t*>(dst_mc),
             (long long)ld_dst_bytes, (long long)rows, (int)(row_bytes / 16));
  ++g_launches;
  return launch_status();
}

// One-off set-up of the peer tensor maps for launches with C_peers == NULL: same tile geometry as the launch
// computes for (C, ldc, M, N, K, e); the upload is synchronous (the maps come from the stack).
extern "C" int pl_gemm_i8_peers_prepare(void* C, void* const* C_peers, int n_peers, void* maps_dev, int ldc, int M,
                                        int N, int K, const pl_epilogue* e) {
  if (!C || !C_peers || !maps_dev || n_peers <= 0 || n_peers > 15 || M <= 0 || N <= 0 || K <= 0 || ldc < N)
    return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (reinterpret_cast<uintptr_t>(maps_dev) & 63) return PL_ERR_ALIGN;
  IgemmParams p{};
  p.BK = pl_conv_block_k(K);
  p.mode = 0;
  rc = common_setup(p, M, N, C, ldc, e, K, 1);
  if (rc) return rc;
  if (e->out_dtype != PL_OUT_INT8 || !p.vec_ok) return PL_ERR_UNSUPPORTED;
  const int PW = p.BN < 128 ? p.BN : 128;
  CUtensorMap maps[15];
  for (int i = 0; i < n_peers; ++i) {
    if (!C_peers[i] || (reinterpret_cast<uintptr_t>(C_peers[i]) & 15)) return PL_ERR_ALIGN;
    rc = encode_2d(&maps[i], C_peers[i], (uint64_t)N, (uint64_t)M, (uint64_t)ldc, PW, kBM);
    if (rc) return rc;
  }
  PL_CUDA_OK(cudaMemcpy(maps_dev, maps, sizeof(CUtensorMap) * n_peers, cudaMemcpyHostToDevice));
  return PL_OK;
}

// ---- allocation-free variant of pl_gemm_i8_peers ----------------------------------------------
// Callers that cannot provide device scratch for the peer tensor maps (CUDA-graph capture set-up, serving
// loops, harnesses that forbid cudaMalloc) use a table that lives in the library image instead: one
// file-scope __device__ array per GPU that loads the library, PL_STATIC_PEERS_MAX maps of 128 bytes.
// The array's NAME is only a host-side stand-in in host code; the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.  All calls on one GPU share the
// table (it keeps the last call's maps): calls that overlap on the GPU -- different streams -- race.
static constexpr int kStaticPeersMax = PL_STATIC_PEERS_MAX;
static __device__ __align__(256) unsigned char g_static_peer_maps[kStaticPeersMax * sizeof(CUtensorMap)];

static int static_peer_maps_address(void** out) {
  constexpr int kMaxDevices = 64;
  static void* addr[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  PL_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return PL_ERR_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* ptr = nullptr;
    PL_CUDA_OK(cudaGetSymbolAddress(&ptr, g_static_peer_maps));  // this device's copy of the array
    if (!ptr || (reinterpret_cast<uintptr_t>(ptr) & 63)) return PL_ERR_ALIGN;
    addr[dev] = ptr;
  }
  *out = addr[dev];
  return PL_OK;
}

extern "C" int pl_gemm_i8_peers_static(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C,
                                       void* const* C_peers, int n_peers, int ldc, int M, int N, int K,
                                       const pl_epilogue* e, pl_stream_t stream) {
  if (n_peers < 0) return PL_ERR_ARG;
  if (n_peers > kStaticPeersMax) {
    fprintf(stderr, "pl_gemm_i8_peers_static: n_peers = %d exceeds the static map table (PL_STATIC_PEERS_MAX = %d); "
                    "use pl_gemm_i8_peers with caller-provided scratch\n", n_peers, kStaticPeersMax);
    return PL_ERR_LIMIT;
  }
  // the timeline trace of the debug builds is the one allocating corner of the launch path
  if (env_int("PL_DBG_TRACE_ON", 0)) return PL_ERR_UNSUPPORTED;
  void* maps = nullptr;
  int rc = static_peer_maps_address(&maps);
  if (rc) return rc;
  return pl_gemm_i8_peers(A, lda, Bt, ldb, C, C_peers, n_peers, maps, ldc, M, N, K, e, stream);
}

extern "C" int pl_last_igemm_plan(int* plan, int n) {
  if (!plan || n <= 0) return 0;
  const int m = n < 10 ? n : 10;
  for (int i = 0; i < m; ++i) plan[i] = g_last_plan[i];
  return m;
}

extern "C" void pl_refresh_knobs(void) {
  KnobCache& k = knobs();
  std::lock_guard<std::mutex> lock(k.mu);
  k.n = 0;
}

extern "C" int pl_gemm_i8(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, int ldc,
                          int M, int N, int K, const pl_epilogue* e, pl_stream_t stream) {
  if (!A || !Bt || !C || M <= 0 || N <= 0 || K <= 0 || lda < K || ldb < K || ldc < N)
    return PL_ERR_ARG;
  int rc = check_epi(e);
  if (rc) return rc;
  if (!driver().ok) return PL_ERR_DRIVER;
  if (driver().cc_major != 10) return PL_ERR_DEVICE;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(Bt) & 15) || (lda % 16) ||
      (ldb % 16))
    return PL_ERR_ALIGN;
  // batch-1 .. batch-64 fc: a weight-streaming kernel (gemv_i8.cu) instead of a mostly empty tensor-core tile
  // (it declines, -100, when its cost model favours the tensor cores)
  if (M <= 64 && env_int("PL_NO_GEMV", 0) == 0) {
    rc = launch_gemv_i8(A, lda, Bt, ldb, C, ldc, M, N, K, e, stream);
    if (rc != -100) return rc;
  }
  IgemmParams p{};
  p.BK = pl_conv_block_k(K);
  p.mode = 0;
  rc = common_setup(p, M, N, C, ldc, e, K, 1);
  if (rc) return rc;
  maybe_enable_mpair(p, e, K, false);
  CUtensorMap ta, tb;
  rc = encode_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, p.BK, kBM << p.mpair);
  if (rc) return rc;
  rc = encode_2d(&tb, Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, p.BK, p.BN);
  if (rc) return rc;
  const BSpec bs{Bt, (uint64_t)K, (uint64_t)N, (uint64_t)ldb};
  return launch_igemm(ta, tb, p, stream, &bs);
}

// same probe with the operand layout of a 32 / 64 / 128-byte K-block (SWIZZLE_32B / 64B / 128B
// descriptors): one iteration issues row_bytes / 32 MMAs of 128 x bn x 32
extern "C" int pl_i8_mma_probe(int ctas, int iters, int bn, int row_bytes, int32_t* out, pl_stream_t stream) {
  if (ctas <= 0 || iters <= 0 || bn < 16 || bn > 256 || (bn % 16) || !out) return PL_ERR_ARG;
  // row_bytes may carry an A start shift in rows in bits 8.. (row_bytes | shift << 8): halo-mode tap starts
  const int shift = row_bytes >> 8;
  row_bytes &= 0xff;
  if (shift < 0 || shift > 31 || (row_bytes != 32 && row_bytes != 64 && row_bytes != 128)) return PL_ERR_ARG;
  const size_t smem = (size_t)(160 + bn) * 128 + 1024;
  PL_CUDA_OK(cudaFuncSetAttribute(i8_mma_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  i8_mma_peak_kernel<<<ctas, 128, smem, (cudaStream_t)stream>>>(iters, bn, out, row_bytes, shift);
  ++g_launches;
  return launch_status();
}

extern "C" int pl_i8_mma_probe_pairs(int ctas, int iters, int bn, int row_bytes, int32_t* out, pl_stream_t stream) {
  if (ctas <= 0 || (ctas & 1) || iters <= 0 || bn < 32 || bn > 256 || (bn % 32) || !out ||
      (row_bytes != 32 && row_bytes != 64 && row_bytes != 128))
    return PL_ERR_ARG;
  const size_t smem = (size_t)(128 + bn / 2) * 128 + 1024;
  PL_CUDA_OK(cudaFuncSetAttribute(i8_mma_peak_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)ctas);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  PL_CUDA_OK(cudaLaunchKernelEx(&cfg, i8_mma_peak_pair_kernel, iters, bn, out, row_bytes));
  ++g_launches;
  return launch_status();
}

extern "C" int pl_i8_mma_peak(int ctas, int iters, int bn, int32_t* out, pl_stream_t stream) {
  if (ctas <= 0 || iters <= 0 || bn < 16 || bn > 256 || (bn % 16) || !out) return PL_ERR_ARG;
  const size_t smem = (size_t)(160 + bn) * 128 + 1024;
  PL_CUDA_OK(cudaFuncSetAttribute(i8_mma_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
  i8_mma_peak_kernel<<<ctas, 128, smem, (cudaStream_t)stream>>>(iters, bn, out, 128, 0);
  ++g_launches;
  return launch_status();
}
