// Channel movement between int8 NHWC ops: concat / split along the channel
// axis, shuffle_channel, relu.  They keep a ShuffleNet-style network int8-NHWC-resident between
// the conv kernels (no calib / layout round trip).  All four are pure HBM streams: every output
// byte is written once from one input byte, so the roofline is 2 bytes of traffic per element.
//
// Reference semantics (NCHW there; a channel-axis memcpy loop becomes a per-pixel byte move here):
//   concat          lite/backends/arm/math/concat.h:29-60, lite/kernels/arm/concat_compute.cc:37-57
//   split           lite/backends/arm/math/split.cc:54-83
//   shuffle_channel lite/backends/arm/math/shuffle_channel.cc:24-76:
//                     out channel (j * group + i) = in channel (i * (C / group) + j)
//   relu            lite/kernels/cuda/relu_compute.cu:23-33 (x >= 0 ? x : 0)
#include "common.cuh"

namespace pl {
namespace {

constexpr int kMaxParts = 8;

struct Parts {
  int8_t* ptr[kMaxParts];
  int ld[kMaxParts];
  int cstart[kMaxParts + 1];  // first wide-tensor channel of part k; cstart[n] = total channels
  int n;
};

template <int V>
struct VecT;
template <>
struct VecT<16> { using type = uint4; };
template <>
struct VecT<4> { using type = uint32_t; };
template <>
struct VecT<1> { using type = uint8_t; };

static inline int grid_for_items(long long items) {
  static const int sms = [] {
    int dev = 0, n = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
  }();
  long long g = (items + 255) / 256;
  const long long cap = (long long)sms * 16;  // grid-stride beyond 16 CTAs (of 256 threads) per SM
  if (g > cap) g = cap;
  return (int)(g < 1 ? 1 : g);
}

// one thread per V-byte chunk of the wide tensor; CONCAT: parts -> wide, else wide -> parts.
// The part of a chunk is found by an unrolled compare chain (no runtime-indexed parameter arrays)
template <int V, bool CONCAT, int NP>
__global__ void __launch_bounds__(256) concat_split_kernel(const Parts parts, int8_t* __restrict__ wide, int ldw,
                                                           const FastDiv dChunks, const uint32_t chunks,
                                                           const uint32_t items) {
  using T = typename VecT<V>::type;
  pdl_launch_dependents();
  pdl_wait();
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < items; i += gridDim.x * 256u) {
    const uint32_t pix = fdiv(i, dChunks);
    const int c = (int)(i - pix * chunks) * V;
    int8_t* pp = parts.ptr[0];
    int ld = parts.ld[0], c0 = 0;
#pragma unroll
    for (int k = 1; k < NP; ++k)  // NP = 2 / 4 / 8 >= parts.n; unused parts start at the total
      if (c >= parts.cstart[k]) {
        pp = parts.ptr[k];
        ld = parts.ld[k];
        c0 = parts.cstart[k];
      }
    T* part = reinterpret_cast<T*>(pp + (long long)pix * ld + (c - c0));
    T* w = reinterpret_cast<T*>(wide + (long long)pix * ldw + c);
    if (CONCAT) *w = *part;
    else *part = *w;
  }
}

// shuffle_channel, group == 2 (ShuffleNet v2): output bytes 2j, 2j+1 = byte j of the two halves.
// W output bytes per thread from W/2 bytes of each half, interleaved with byte permutes.  The halves
// are two row pointers: the same tensor (plain shuffle_channel) or two tensors (concat + shuffle fused)
template <int W>
__global__ void __launch_bounds__(256) shuffle2_kernel(const int8_t* __restrict__ xa, int lda,
                                                       const int8_t* __restrict__ xb, int ldb, int8_t* __restrict__ y,
                                                       int ldy, const FastDiv dChunks, const uint32_t chunks,
                                                       const uint32_t items) {
  pdl_launch_dependents();
  pdl_wait();
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < items; i += gridDim.x * 256u) {
    const uint32_t pix = fdiv(i, dChunks);
    const int ch = (int)(i - pix * chunks);
    const int8_t* ra = xa + (long long)pix * lda + ch * (W / 2);
    const int8_t* rb = xb + (long long)pix * ldb + ch * (W / 2);
    int8_t* out = y + (long long)pix * ldy + ch * W;
    if (W == 16) {
      const uint2 a = *reinterpret_cast<const uint2*>(ra);
      const uint2 b = *reinterpret_cast<const uint2*>(rb);
      uint4 o;
      o.x = __byte_perm(a.x, b.x, 0x5140);
      o.y = __byte_perm(a.x, b.x, 0x7362);
      o.z = __byte_perm(a.y, b.y, 0x5140);
      o.w = __byte_perm(a.y, b.y, 0x7362);
      *reinterpret_cast<uint4*>(out) = o;
    } else if (W == 4) {
      const uint32_t a = *reinterpret_cast<const uint16_t*>(ra);
      const uint32_t b = *reinterpret_cast<const uint16_t*>(rb);
      *reinterpret_cast<uint32_t*>(out) = __byte_perm(a, b, 0x5140);
    } else {  // W == 2: no alignment at all
      out[0] = ra[0];
      out[1] = rb[0];
    }
  }
}

// any group: one thread per output byte (W == 1) or per aligned output word gathered bytewise (W == 4)
template <int W>
__global__ void __launch_bounds__(256) shuffle_any_kernel(const int8_t* __restrict__ x, int ldx,
                                                          int8_t* __restrict__ y, int ldy, int cg,
                                                          const FastDiv dGroup, int group, const FastDiv dChunks,
                                                          const uint32_t chunks, const uint32_t items) {
  pdl_launch_dependents();
  pdl_wait();
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < items; i += gridDim.x * 256u) {
    const uint32_t pix = fdiv(i, dChunks);
    const int co = (int)(i - pix * chunks) * W;
    const int8_t* row = x + (long long)pix * ldx;
    uint32_t word = 0;
#pragma unroll
    for (int t = 0; t < W; ++t) {
      const uint32_t j = fdiv((uint32_t)(co + t), dGroup);  // co = j * group + i
      const int ig = co + t - (int)j * group;
      word |= (uint32_t)(uint8_t)row[ig * cg + (int)j] << (8 * t);
    }
    int8_t* out = y + (long long)pix * ldy + co;
    if (W == 4) *reinterpret_cast<uint32_t*>(out) = word;
    else *out = (int8_t)word;
  }
}

__global__ void __launch_bounds__(256) relu_i8_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ y,
                                                      long long n, int hi, int vec) {
  pdl_launch_dependents();
  pdl_wait();
  const long long stride = (long long)gridDim.x * 256;
  const uint32_t hi4 = (uint32_t)(hi & 0xff) * 0x01010101u;
  if (vec) {
    const long long nv = n >> 4;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nv; i += stride) {
      uint4 v = reinterpret_cast<const uint4*>(x)[i];
      v.x = __vmins4(__vmaxs4(v.x, 0u), hi4);
      v.y = __vmins4(__vmaxs4(v.y, 0u), hi4);
      v.z = __vmins4(__vmaxs4(v.z, 0u), hi4);
      v.w = __vmins4(__vmaxs4(v.w, 0u), hi4);
      reinterpret_cast<uint4*>(y)[i] = v;
    }
    for (long long i = (nv << 4) + blockIdx.x * 256ll + threadIdx.x; i < n; i += stride) {
      const int v = x[i] > 0 ? x[i] : 0;
      y[i] = (int8_t)(v > hi ? hi : v);
    }
  } else {
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += stride) {
      const int v = x[i] > 0 ? x[i] : 0;
      y[i] = (int8_t)(v > hi ? hi : v);
    }
  }
}

inline bool aligned(const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) & (uintptr_t)(a - 1)) == 0; }

// widest chunk (16 / 4 / 1 bytes) every part boundary, pitch and base pointer is a multiple of
int chunk_bytes(const Parts& parts, const void* wide, int ldw) {
  int v = 16;
  while (v > 1) {
    bool ok = aligned(wide, v) && ldw % v == 0;
    for (int k = 0; k < parts.n && ok; ++k)
      ok = aligned(parts.ptr[k], v) && parts.ld[k] % v == 0 && parts.cstart[k + 1] % v == 0;
    if (ok) break;
    v = v == 16 ? 4 : 1;
  }
  return v;
}

template <bool CONCAT>
int run_concat_split(Parts& parts, int8_t* wide, int ldw, int64_t pixels, pl_stream_t stream) {
  const int ctot = parts.cstart[parts.n];
  const int v = chunk_bytes(parts, wide, ldw);
  const uint32_t chunks = (uint32_t)(ctot / v);
  const FastDiv d = make_fastdiv(chunks);
  // 32-bit item indices: long tensors go out in slabs of whole pixels
  const int64_t slab = ((1ll << 31) - 1) / chunks;
  for (int64_t p0 = 0; p0 < pixels; p0 += slab) {
    const int64_t np = pixels - p0 < slab ? pixels - p0 : slab;
    Parts q = parts;
    for (int k = 0; k < parts.n; ++k) q.ptr[k] = parts.ptr[k] + p0 * parts.ld[k];
    int8_t* w = wide + p0 * ldw;
    const uint32_t items = (uint32_t)(np * chunks);
    const int grid = grid_for_items(items);
#define PL_CS_LAUNCH(V_, NP_) \
  launch_pdl(concat_split_kernel<V_, CONCAT, NP_>, grid, 256, 0, (cudaStream_t)stream, q, w, ldw, d, chunks, items)
#define PL_CS_PARTS(V_)                      \
  do {                                       \
    if (parts.n <= 2) PL_CS_LAUNCH(V_, 2);   \
    else if (parts.n <= 4) PL_CS_LAUNCH(V_, 4); \
    else PL_CS_LAUNCH(V_, 8);                \
  } while (0)
    if (v == 16) PL_CS_PARTS(16);
    else if (v == 4) PL_CS_PARTS(4);
    else PL_CS_PARTS(1);
#undef PL_CS_PARTS
#undef PL_CS_LAUNCH
    ++g_launches;
  }
  return launch_status();
}

int fill_parts(Parts& parts, int n, int8_t* const* ptrs, const int* cs, const int* lds) {
  if (n < 1 || n > kMaxParts || !ptrs || !cs || !lds) return PL_ERR_ARG;
  parts.n = n;
  parts.cstart[0] = 0;
  for (int k = 0; k < n; ++k) {
    if (!ptrs[k] || cs[k] <= 0 || lds[k] < cs[k]) return PL_ERR_ARG;
    parts.ptr[k] = ptrs[k];
    parts.ld[k] = lds[k];
    parts.cstart[k + 1] = parts.cstart[k] + cs[k];
  }
  for (int k = n; k < kMaxParts; ++k) {
    parts.ptr[k] = nullptr;
    parts.ld[k] = 0;
    parts.cstart[k + 1] = parts.cstart[n];
  }
  return PL_OK;
}

}  // namespace
}  // namespace pl

using namespace pl;

extern "C" int pl_concat_i8_nhwc(int n_in, const int8_t* const* xs, const int* cs, const int* ldxs, int8_t* y,
                                 int ldy, int64_t pixels, pl_stream_t stream) {
  Parts parts;
  const int rc = fill_parts(parts, n_in, const_cast<int8_t* const*>(reinterpret_cast<const int8_t* const*>(xs)), cs, ldxs);
  if (rc) return rc;
  if (!y || pixels < 0 || ldy < parts.cstart[n_in]) return PL_ERR_ARG;
  if (pixels == 0) return PL_OK;
  return run_concat_split<true>(parts, y, ldy, pixels, stream);
}

extern "C" int pl_split_i8_nhwc(const int8_t* x, int ldx, int n_out, int8_t* const* ys, const int* cs,
                                const int* ldys, int64_t pixels, pl_stream_t stream) {
  Parts parts;
  const int rc = fill_parts(parts, n_out, ys, cs, ldys);
  if (rc) return rc;
  if (!x || pixels < 0 || ldx < parts.cstart[n_out]) return PL_ERR_ARG;
  if (pixels == 0) return PL_OK;
  return run_concat_split<false>(parts, const_cast<int8_t*>(x), ldx, pixels, stream);
}

extern "C" int pl_shuffle_channel_i8_nhwc(const int8_t* x, int ldx, int8_t* y, int ldy, int C, int group,
                                          int64_t pixels, pl_stream_t stream) {
  if (!x || !y || C <= 0 || group <= 0 || C % group || ldx < C || ldy < C || pixels < 0 || x == y) return PL_ERR_ARG;
  if (pixels == 0) return PL_OK;
  const int cg = C / group;
  // widest output chunk the shapes allow
  int W = 1;
  bool two = group == 2;
  if (two && cg % 8 == 0 && aligned(x, 8) && ldx % 8 == 0 && aligned(y, 16) && ldy % 16 == 0) W = 16;
  else if (two && cg % 2 == 0 && aligned(x, 2) && ldx % 2 == 0 && aligned(y, 4) && ldy % 4 == 0) W = 4;
  else {
    two = false;
    if (C % 4 == 0 && aligned(y, 4) && ldy % 4 == 0) W = 4;
  }
  const uint32_t chunks = (uint32_t)(C / W);
  const FastDiv d = make_fastdiv(chunks), dg = make_fastdiv((uint32_t)group);
  const int64_t slab = ((1ll << 31) - 1) / chunks;
  for (int64_t p0 = 0; p0 < pixels; p0 += slab) {
    const int64_t np = pixels - p0 < slab ? pixels - p0 : slab;
    const int8_t* xs = x + p0 * ldx;
    int8_t* ys = y + p0 * ldy;
    const uint32_t items = (uint32_t)(np * chunks);
    const int grid = grid_for_items(items);
    if (two && W == 16) launch_pdl(shuffle2_kernel<16>, grid, 256, 0, (cudaStream_t)stream, xs, ldx, xs + cg, ldx, ys, ldy, d, chunks, items);
    else if (two) launch_pdl(shuffle2_kernel<4>, grid, 256, 0, (cudaStream_t)stream, xs, ldx, xs + cg, ldx, ys, ldy, d, chunks, items);
    else if (W == 4) launch_pdl(shuffle_any_kernel<4>, grid, 256, 0, (cudaStream_t)stream, xs, ldx, ys, ldy, cg, dg, group, d, chunks, items);
    else launch_pdl(shuffle_any_kernel<1>, grid, 256, 0, (cudaStream_t)stream, xs, ldx, ys, ldy, cg, dg, group, d, chunks, items);
    ++g_launches;
  }
  return launch_status();
}

// concat(a, b) along channels followed by shuffle_channel(group 2) in one pass: the tail of every
// ShuffleNet-v2 unit.  out[2j] = a[j], out[2j + 1] = b[j]; 2 bytes of traffic per element instead of 4
extern "C" int pl_concat_shuffle2_i8_nhwc(const int8_t* a, int lda, const int8_t* b, int ldb, int8_t* y, int ldy,
                                          int c_half, int64_t pixels, pl_stream_t stream) {
  if (!a || !b || !y || c_half <= 0 || lda < c_half || ldb < c_half || ldy < 2 * c_half || pixels < 0) return PL_ERR_ARG;
  if (pixels == 0) return PL_OK;
  int W = 2;
  if (c_half % 8 == 0 && aligned(a, 8) && aligned(b, 8) && lda % 8 == 0 && ldb % 8 == 0 && aligned(y, 16) && ldy % 16 == 0)
    W = 16;
  else if (c_half % 2 == 0 && aligned(a, 2) && aligned(b, 2) && lda % 2 == 0 && ldb % 2 == 0 && aligned(y, 4) &&
           ldy % 4 == 0)
    W = 4;
  const uint32_t chunks = (uint32_t)(2 * c_half / W);
  const FastDiv d = make_fastdiv(chunks);
  const int64_t slab = ((1ll << 31) - 1) / chunks;
  for (int64_t p0 = 0; p0 < pixels; p0 += slab) {
    const int64_t np = pixels - p0 < slab ? pixels - p0 : slab;
    const uint32_t items = (uint32_t)(np * chunks);
    const int grid = grid_for_items(items);
    const int8_t* as = a + p0 * lda;
    const int8_t* bs = b + p0 * ldb;
    int8_t* ys = y + p0 * ldy;
    if (W == 16) launch_pdl(shuffle2_kernel<16>, grid, 256, 0, (cudaStream_t)stream, as, lda, bs, ldb, ys, ldy, d, chunks, items);
    else if (W == 4) launch_pdl(shuffle2_kernel<4>, grid, 256, 0, (cudaStream_t)stream, as, lda, bs, ldb, ys, ldy, d, chunks, items);
    else launch_pdl(shuffle2_kernel<2>, grid, 256, 0, (cudaStream_t)stream, as, lda, bs, ldb, ys, ldy, d, chunks, items);
    ++g_launches;
  }
  return launch_status();
}

extern "C" int pl_relu_i8(const int8_t* x, int8_t* y, int64_t n, int hi, pl_stream_t stream) {
  if (!x || !y || n < 0 || hi < 0 || hi > 127) return PL_ERR_ARG;
  if (n == 0) return PL_OK;
  const int vec = aligned(x, 16) && aligned(y, 16);
  const int grid = grid_for_items(vec ? (n + 15) / 16 : n);
  launch_pdl(relu_i8_kernel, grid, 256, 0, (cudaStream_t)stream, x, y, (long long)n, hi, vec);
  ++g_launches;
  return launch_status();
}
