// gemv_i8.cu -- skinny int8 GEMM (batch-1 .. batch-64 fc / mul / matmul) as a bandwidth kernel: 8 rows of A per
// CTA row (blockIdx.y), the weights re-read from L2 by every row group.
//
// Replaces the role of lite/backends/arm/math/gemv_arm_int8.cc:25-143 (gemv_int8: one output per weight
// row, K split over NEON lanes, write_gemv_out epilogue).  With M <= 8 the weights Bt[N, K] are read once and
// nothing is reused: the tensor-core tile (128 rows of A) would run 1 / 128 full and its 32 k-blocks of ~450
// cycles per CTA bound 1 x 4096 x 4096 at 18 us (0.9 TB/s).  Here a warp owns 4 consecutive weight rows, its
// lanes stride K in 16-byte vectors (4 independent 16-byte streaming loads per lane and step, the A chunk is
// loaded once per step and reused by the 4 rows), dp4a accumulates in int32 (exact), a butterfly reduces the
// lanes, and -- when N alone cannot fill the machine -- K is split over the warps of the CTA and reduced
// through shared memory (no global scratch, no atomics).  Same fused epilogue as the GEMM.
#include <cmath>

#include "common.cuh"

namespace pl {

struct GemvParams {
  const int8_t* A;
  const int8_t* Bt;
  void* C;
  int lda, ldb, ldc;
  int M, N, K;
  int ks;  // k-splits per row block (1, 2, 4, 8): warps of one CTA
  EpiParams epi;
};

static constexpr int kGemvWarps = 8;
static constexpr int kGemvRows = 4;  // weight rows per warp

template <int MT>
__global__ void __launch_bounds__(32 * kGemvWarps) gemv_i8_kernel(const GemvParams p) {
  __shared__ int red[kGemvWarps][MT * kGemvRows];
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int upc = kGemvWarps / p.ks;  // row blocks per CTA
  const int unit = blockIdx.x * upc + warp / p.ks;
  const int ks = warp % p.ks;
  const int n0 = unit * kGemvRows;
  const int chunks = (p.K + 15) >> 4;
  const int cps = (chunks + p.ks - 1) / p.ks;
  const int c_begin = ks * cps;
  const int c_end = min(chunks, c_begin + cps);
  const int tail = p.K & 15;  // bytes of the last chunk that belong to K (0 = all 16)
  const int m0 = (int)blockIdx.y * 8;  // this CTA row's 8 rows of A / C (MT == 8 whenever gridDim.y > 1)
  const int Mloc = min(p.M - m0, MT);
  int acc[MT][kGemvRows];
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) acc[m][r] = 0;
  pdl_wait();  // A comes from the previous kernel
  if (n0 < p.N) {
    const int8_t* brow[kGemvRows];
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) brow[r] = p.Bt + (size_t)min(n0 + r, p.N - 1) * p.ldb;
#pragma unroll 2
    for (int c = c_begin + lane; c < c_end; c += 32) {
      int4 b[kGemvRows];
#pragma unroll
      for (int r = 0; r < kGemvRows; ++r) b[r] = __ldcs(reinterpret_cast<const int4*>(brow[r]) + c);
      if (tail && c == chunks - 1) {
        // the pad bytes of A's last chunk are the caller's: zero the weights beyond K instead
#pragma unroll
        for (int r = 0; r < kGemvRows; ++r) {
          int w[4] = {b[r].x, b[r].y, b[r].z, b[r].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int keep = tail - 4 * j;  // bytes of word j inside K
            w[j] = keep >= 4 ? w[j] : (keep <= 0 ? 0 : (w[j] & ((1 << (8 * keep)) - 1)));
          }
          b[r] = make_int4(w[0], w[1], w[2], w[3]);
        }
      }
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        if (m < Mloc) {
          const int4 a = __ldg(reinterpret_cast<const int4*>(p.A + (size_t)(m0 + m) * p.lda) + c);
#pragma unroll
          for (int r = 0; r < kGemvRows; ++r) {
            int s = acc[m][r];
            s = __dp4a(a.x, b[r].x, s);
            s = __dp4a(a.y, b[r].y, s);
            s = __dp4a(a.z, b[r].z, s);
            s = __dp4a(a.w, b[r].w, s);
            acc[m][r] = s;
          }
        }
      }
    }
  }
#pragma unroll
  for (int m = 0; m < MT; ++m)
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) {
      int v = acc[m][r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][m * kGemvRows + r] = v;
    }
  __syncthreads();
  if (ks != 0 || n0 >= p.N || lane >= MT * kGemvRows) return;
  int sum = 0;
  for (int s = 0; s < p.ks; ++s) sum += red[warp + s][lane];
  const int m = lane / kGemvRows, n = n0 + lane % kGemvRows;
  if (m >= Mloc || n >= p.N) return;
  const EpiParams& e = p.epi;
  const size_t o = (size_t)(m0 + m) * p.ldc + n;
  if (e.out_dtype == PL_OUT_INT32) {
    reinterpret_cast<int*>(p.C)[o] = sum;
    return;
  }
  const float v = epi_value(sum, __ldg(e.scale + n), e.bias ? __ldg(e.bias + n) : 0.f, e);
  if (e.out_dtype == PL_OUT_FP32)
    reinterpret_cast<float*>(p.C)[o] = v;
  else
    reinterpret_cast<int8_t*>(p.C)[o] = (int8_t)round_sat_i8(v, e.round_mode);
}

// M <= 64, no fused residual: returns PL_OK after launching, or -100 when the shape is not for this kernel
int launch_gemv_i8(const int8_t* A, int lda, const int8_t* Bt, int ldb, void* C, int ldc, int M, int N, int K,
                   const pl_epilogue* e, pl_stream_t stream) {
  if (M > 64 || e->residual) return -100;
  {
    // measured on B200 (tools/gemv_time.py): this kernel streams weights at 4.1-5.9 TB/s for M = 1 but is
    // dp4a-bound at ~13 GMAC/us as M grows (8 x 16384^2: 137 us against 57 us on the tensor cores, whose skinny
    // launches have a ~17 us floor and stream at ~4.7 TB/s); take whichever the model predicts faster
    const double nk = (double)N * K;
    const double t_gemv = fmax(nk / 5.5e6, (double)M * nk / 13e6);
    // (small tensor-core GEMMs: 10.7 us for 64 x 1000 x 1024 in isolation, 16 us inside MobileNetV1)
    const double t_tc = fmax(M <= 8 ? 17.0 : 11.0, nk / 4.7e6);
    if (t_gemv > t_tc) return -100;
  }
  GemvParams p{};
  p.A = A; p.Bt = Bt; p.C = C; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.M = M; p.N = N; p.K = K;
  p.epi = make_epi(e, N);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int units = (N + kGemvRows - 1) / kGemvRows;
  const int chunks = (K + 15) / 16;
  // split K over the warps of a CTA while the row blocks alone leave warps idle and every split still
  // streams at least 1 KB (64 chunks) per weight row
  const unsigned gy = (unsigned)((M + 7) / 8);
  int ks = 1;
  while (ks < kGemvWarps && (long long)units * ks * gy < 2ll * sms * kGemvWarps && chunks / (ks * 2) >= 64) ks *= 2;
  p.ks = ks;
  const int upc = kGemvWarps / ks;
  const unsigned grid = (unsigned)((units + upc - 1) / upc);
  cudaStream_t st = (cudaStream_t)stream;
  if (M <= 1) launch_pdl(gemv_i8_kernel<1>, dim3(grid), dim3(32 * kGemvWarps), 0, st, p);
  else if (M <= 2) launch_pdl(gemv_i8_kernel<2>, dim3(grid), dim3(32 * kGemvWarps), 0, st, p);
  else if (M <= 4) launch_pdl(gemv_i8_kernel<4>, dim3(grid), dim3(32 * kGemvWarps), 0, st, p);
  else launch_pdl(gemv_i8_kernel<8>, dim3(grid, gy), dim3(32 * kGemvWarps), 0, st, p);
  ++g_launches;
  return launch_status();
}

}  // namespace pl
