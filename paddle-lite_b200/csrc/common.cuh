// common.cuh -- sm_100a PTX wrappers (mbarrier, TMA, tcgen05/TMEM) and the
// scalar epilogue shared by every kernel of the int8 hot path.
//
// Epilogue semantics follow the reference (file:line in chenjiaoAngel/Paddle-Lite):
//   v = fma((float)acc, scale, bias)      lite/backends/arm/math/gemm_prepacked_int8.cc:664-700
//   relu / relu6 / leaky                  lite/backends/arm/math/gemv_arm_int8.cc:111-116
//   int8: round-half-away, clamp +-127    lite/backends/arm/math/gemv_arm_int8.cc:122-124,
//                                         lite/backends/arm/math/saturate.h:130-147
//   legacy: clamp [-128,127] + rint       lite/backends/cuda/math/utils.h:61-66
// Built with -fmad=false: the only fused multiply-adds are the explicit
// __fmaf_rn below, so the CPU oracle (fmaf, -ffp-contract=off) matches bit for bit.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pl_int8.h"

namespace pl {

extern std::atomic<unsigned long long> g_launches;  // host-side launch counter (capi.cu)

#define PL_CUDA_OK(expr)                         \
  do {                                           \
    cudaError_t _e = (expr);                     \
    if (_e != cudaSuccess) return (int)_e;       \
  } while (0)

static inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? PL_OK : (int)e;
}

// ----------------------------------------------------------------------------
// programmatic dependent launch: consecutive layers of a network are launched with
// programmaticStreamSerialization so that the next kernel's CTAs become resident (barrier
// init, TMEM allocation, descriptor prefetch, scale/bias staging) while the tail of the
// previous kernel drains.  Every kernel calls pdl_launch_dependents() at its top and
// pdl_wait() before it first touches memory another kernel may have produced or may still
// be reading (the wait returns once the whole previous grid has completed and flushed).
// PL_PDL=0 launches everything fully serialised (A/B switch).
// ----------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // capi.cu

template <typename... KArgs, typename... Args>
static inline void launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      int cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (cluster_x > 1) {  // thread-block clusters (CTA pairs: the two CTAs land on the two SMs of one TPC)
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = (unsigned)cluster_x;
    at[n].val.clusterDim.y = 1;
    at[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = (unsigned)n;
  cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);  // errors surface through launch_status()
}
template <typename... KArgs, typename... Args>
static inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  launch_pdl_cluster(kernel, grid, block, smem, st, 1, static_cast<Args&&>(args)...);
}

// ----------------------------------------------------------------------------
// epilogue math
// ----------------------------------------------------------------------------
__device__ __forceinline__ float apply_act(float v, int act, float arg) {
  switch (act) {
    case PL_ACT_RELU: return v > 0.f ? v : 0.f;
    case PL_ACT_RELU6: return v > 0.f ? (v > arg ? arg : v) : 0.f;
    case PL_ACT_LEAKY: return v > 0.f ? v : __fmul_rn(v, arg);
    default: return v;
  }
}

// CUDA-legacy activation (activation.cu:91-181): vec4 path fmaxf(v*alpha, v)
// when the channel count is a multiple of 4, else v>0 ? v : v*alpha.
__device__ __forceinline__ float apply_act_legacy(float v, int act, float arg, bool vec4) {
  if (act == PL_ACT_RELU || act == PL_ACT_LEAKY) {
    const float alpha = act == PL_ACT_RELU ? 0.f : arg;
    return vec4 ? fmaxf(__fmul_rn(v, alpha), v) : (v > 0.f ? v : __fmul_rn(v, alpha));
  }
  if (act == PL_ACT_RELU6) return v > 0.f ? (v > arg ? arg : v) : 0.f;
  return v;
}

// round + saturate to int8, returned as int in [-128,127].  No FRND/F2I: the
// conversion pipe is 4-8x narrower than the FMA pipe and the epilogue runs once
// per output byte, so both conventions are built from FADDs with explicit
// rounding modes (bit-exact with roundf()/rintf(), see tests/test_oracle.py):
//   ARM    : m = floor(RZ(min(|v|,127) + 0.5)) via RD(t + 2^23); sign re-applied
//   legacy : low byte of RNE(clamp(v,-128,127) + 1.5*2^23)
__device__ __forceinline__ int round_sat_i8(float v, int round_mode) {
  if (round_mode == PL_ROUND_CUDA_LEGACY) {
    const float c = fminf(fmaxf(v, -128.f), 127.f);
    const int bits = __float_as_int(__fadd_rn(c, 12582912.f));
    return (int)(int8_t)(bits & 0xff);
  }
  const float a = fminf(fabsf(v), 127.f);
  const float t = __fadd_rz(a, 0.5f);
  const int m = __float_as_int(__fadd_rd(t, 8388608.f)) & 0xff;
  return v < 0.f ? -m : m;
}

// exact int -> float for |acc| < 2^22 without the conversion pipe
__device__ __forceinline__ float small_int2float(int acc) {
  return __fadd_rn(__int_as_float(acc + 0x4B400000), -12582912.f);
}

// division by a launch-time constant: Granlund-Montgomery round-up magic
struct FastDiv {  // q = (n * mul) >> (32 + sh), exact for n < 2^31 (round-up magic)
  uint32_t mul, sh, d;
};
static inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  if (d == 1) {
    f.mul = 0;
    f.sh = 0;
    return f;
  }
  uint32_t l = 0;
  while ((1u << l) < d) ++l;  // ceil(log2 d)
  const uint64_t m = ((1ull << (32 + l)) + d - 1) / d;  // 33-bit magic, top bit implicit
  f.mul = (uint32_t)(m - (1ull << 32));
  f.sh = l;
  return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv f) {
  if (f.d == 1) return n;
  const uint32_t t = __umulhi(n, f.mul);
  return (t + ((n - t) >> 1)) >> (f.sh - 1);
}

struct EpiParams {
  const float* scale;
  const float* bias;
  const int8_t* residual;
  int ld_res;
  float res_scale;
  int act;
  float act_arg;
  int out_dtype;
  int round_mode;
  int vec4;  // legacy activation flavour (n_out % 4 == 0)
};

static inline EpiParams make_epi(const pl_epilogue* e, int n_out) {
  EpiParams p;
  p.scale = e->scale;
  p.bias = e->bias;
  p.residual = e->residual;
  p.ld_res = e->ld_res;
  p.res_scale = e->res_scale;
  p.act = e->act;
  p.act_arg = e->act_arg;
  p.out_dtype = e->out_dtype;
  p.round_mode = e->round_mode;
  p.vec4 = (n_out % 4 == 0);
  return p;
}

// one accumulator -> pre-rounding float
__device__ __forceinline__ float epi_value(int acc, float s, float b, const EpiParams& p) {
  float v = __fmaf_rn(__int2float_rn(acc), s, b);
  if (p.round_mode == PL_ROUND_CUDA_LEGACY) return apply_act_legacy(v, p.act, p.act_arg, p.vec4);
  return apply_act(v, p.act, p.act_arg);
}

__device__ __forceinline__ uint32_t pack4_i8(int a, int b, int c, int d) {
  return (uint32_t)(a & 0xff) | ((uint32_t)(b & 0xff) << 8) | ((uint32_t)(c & 0xff) << 16) |
         ((uint32_t)(d & 0xff) << 24);
}

// ----------------------------------------------------------------------------
// mbarrier
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// one lane of the (converged) warp; the compiler then keeps single-issue instructions
// (TMA, tcgen05.mma / commit) and their operands on the uniform datapath
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must end in a trap (launch failure reported to
// the caller), never in a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) __trap();
  }
}

// 4-byte asynchronous global -> shared copy (LDGSTS); src_bytes < 4 zero-fills the rest
__device__ __forceinline__ void cp_async_4(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_8(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// this thread's arrival on `bar` happens when all its earlier cp.async copies have landed
// (.noinc: the arrival is one of the barrier's expected arrivals)
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// pure polling variant (mbarrier.test_wait never suspends the thread)
__device__ __forceinline__ void mbar_wait_poll(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  long long t0 = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!ok) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000LL) __trap();
    }
  } while (!ok);
}

// ----------------------------------------------------------------------------
// TMA
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const void* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(desc) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* desc, uint32_t bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// L2 eviction-priority hints for loads of bytes that are read exactly once (a fused residual, the A rows of a
// launch with one N tile): they leave the cache first, so the tensor being WRITTEN stays for the next layer
// (the encodings createpolicy.fractional.L2::evict_first / evict_normal .b64 %0, 1.0 return: as constants they can
// live on the uniform datapath of a single-thread role loop instead of costing two R2UR per TMA instruction)
static constexpr uint64_t kL2EvictFirst = 0x12F0000000000000ull, kL2EvictNormal = 0x16F0000000000000ull;
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_hint(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1,
                                                 int c2, int c3, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d_hint(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1, int c2,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}
// 16-byte read-only load with an L2 eviction hint
__device__ __forceinline__ uint4 ldg_v4_hint(const void* p, uint64_t policy) {
  uint4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p), "l"(policy));
  return v;
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// 1-D bulk copy global -> shared (no tensor map): `bytes` contiguous bytes, 16-byte aligned both sides
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tma_store_4d(const void* desc, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(desc),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// smem -> global tiled store (bulk async-group completion); rows/cols outside the
// tensor map's extent are clipped by the hardware
__device__ __forceinline__ void tma_store_2d(const void* desc, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(desc),
               "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all but the N most recent bulk groups have finished READING their smem source
template <int N>
__device__ __forceinline__ void bulk_wait_group_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_group() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// im2col-mode load of `pixelsPerColumn` output pixels starting at base pixel
// (w,h,n), channels [c, c+channelsPerPixel), filter offset (off_w, off_h).
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t dst, const void* desc, uint32_t bar,
                                                   int c, int w, int h, int n, uint16_t off_w,
                                                   uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}

__device__ __forceinline__ void tma_load_im2col_4d_hint(uint32_t dst, const void* desc, uint32_t bar,
                                                        int c, int w, int h, int n, uint16_t off_w,
                                                        uint16_t off_h, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8}, %9;" ::"r"(dst),
      "l"(desc), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h), "l"(policy)
      : "memory");
}

// ----------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, int8 x int8 -> int32, issued by ONE thread
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                        uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued MMAs of this thread are done
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   bar)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 16 consecutive 32-bit columns: thread t gets row (lane base + t)
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// 32 lanes x 32 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

// ----------------------------------------------------------------------------
// CTA pairs (cta_group::2): two CTAs of one cluster on the two SMs of a TPC run ONE MMA of
// M = 256: each CTA holds its own 128 rows of A, HALF of the B tile (N / 2 rows) and the
// 128 x N accumulator block of its rows in its own TMEM.  The leader (cluster rank 0) issues
// the MMAs; both CTAs issue TMA loads into their own smem that complete on the LEADER's
// mbarrier; tcgen05.commit multicasts its arrival to the same barrier offset in both CTAs.
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {  // every thread of every CTA of the cluster
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// arrival on a barrier of another CTA of the cluster (shared::cluster address)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const void* desc, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(desc), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(uint32_t dst, const void* desc, uint32_t bar_cluster, int c0, int c1,
                                                int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(desc), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col_4d_2sm(uint32_t dst, const void* desc, uint32_t bar_cluster, int c,
                                                       int w, int h, int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
      "l"(desc), "r"(bar_cluster), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t smem_dst, uint32_t ncols) {  // one warp of EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] (+)= A (128 rows per CTA) * B^T (N / 2 rows per CTA), issued by ONE thread of the leader
__device__ __forceinline__ void umma_i8_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrival on the barrier at this offset in BOTH CTAs of the pair once the pair's MMAs issued so far are done
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}

// one 16-byte store through an NVSwitch multicast address: the switch replicates it into the same offset of every
// GPU bound to the multicast object (NVLS).  The .f32 type only names the vector layout: the bits are copied.
__device__ __forceinline__ void multimem_st_v4(void* mc_addr, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// SM100 shared-memory matrix descriptor for a K-major operand tile whose rows
// are `row_bytes` (32 / 64 / 128) wide and swizzled with the matching TMA
// swizzle mode.  Fields (cute/arch/mma_sm100_desc.hpp): start>>4 [0,14),
// LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout [61,64).
__host__ __device__ __forceinline__ uint64_t make_smem_desc_base(uint32_t row_bytes) {
  const uint64_t layout = row_bytes == 128 ? 2ull : (row_bytes == 64 ? 4ull : 6ull);
  const uint64_t sbo = (8ull * row_bytes) >> 4;  // 8-row core-matrix group pitch
  return (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}
// Un-swizzled K-major operand whose rows are 16 bytes apart: a core matrix is 8 rows x 16 bytes
// (128 contiguous bytes), SBO = 128 B between 8-row groups, and LBO = 16 B between the two 16-byte
// halves of a K = 32 step -- i.e. the second half of row r IS the first half of row r + 1.  Over a
// plain array of 16-byte pixels this reads A[r][0..31] = pixels r, r + 1: two horizontally adjacent
// filter taps of the space-to-depth stem per MMA, with no im2col copy at all.
__host__ __device__ __forceinline__ uint64_t make_smem_desc_pix16() {
  return (1ull << 16) | (8ull << 32) | (1ull << 46);
}
// instruction descriptor: D=S32, A=B=signed int8, K-major both, M=128 (256: the two CTAs of a pair)
__host__ __device__ __forceinline__ uint32_t make_idesc_i8(uint32_t n, uint32_t m = 128u) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

// scheduling hint of pl_set_tile_order (igemm_i8.cu): should the next launch of this thread walk its tiles last-to-first?
bool next_tile_order_reversed();

}  // namespace pl
