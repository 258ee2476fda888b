// What does the MMA-issuing thread pay per tile?  One thread issues batches of `nb` tcgen05.mma (128 x bn x 32, resident
// operands) into alternating accumulators; between batches it runs the synchronising instructions of the real tile loop,
// selected by `mode` bits.  Prints cycles per batch (clock64 of the issuing thread, CTA 0) and the share of the time the
// thread spent inside the MMA issue loop.
//   1  one tcgen05.commit after the batch        2  a second commit
//   4  one mbarrier.try_wait (already complete)  8  a second try_wait
//  16  tcgen05.fence::after_thread_sync after the waits
//  32  waits BEFORE the commits (software-pipelined order) instead of after
//  64  every batch accumulates onto its accumulator (no fresh start)
// 128  no clock reads inside the loop
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I paddle-lite_b200/csrc -I include tools/sync_probe.cu -o tools/bin/sync_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "common.cuh"

using namespace pl;

template <int NB>
__global__ void __launch_bounds__(128, 1) sync_probe_kernel(int iters, int bn, int mode, long long* out, int a_pix16) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[4];
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sa = base, sb = base + 160 * 128;
  for (uint32_t i = threadIdx.x; i < (160u + (uint32_t)bn) * 128u / 4u; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x01010101u));
  fence_proxy_async();
  const int warp = threadIdx.x >> 5;
  const uint32_t b_commit = smem_u32(&bars[0]), b_commit2 = smem_u32(&bars[1]), b_done = smem_u32(&bars[2]),
                 b_ready = smem_u32(&bars[3]);
  if (threadIdx.x == 0) {
    mbar_init(b_commit, 1);
    mbar_init(b_commit2, 1);
    mbar_init(b_done, 1);
    mbar_init(b_ready, 1);
    fence_barrier_init();
    mbar_arrive(b_ready);  // phase 0 of b_ready is complete for the whole run: try_wait(parity 0) succeeds at once
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
    const uint64_t hi = make_smem_desc_base(128u);
    // a_pix16: the space-to-depth stems' A layout (no swizzle, rows 16 bytes apart, K chunks overlap the next row)
    const uint64_t da = (a_pix16 ? make_smem_desc_pix16() : hi) | (uint64_t)((sa >> 4) & 0x3fff), db = hi | (uint64_t)((sb >> 4) & 0x3fff);
    long long t_issue = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tmem + (uint32_t)((i & 1) * bn);
      const long long a0 = (mode & 128) ? 0 : clock64();
      const uint32_t keep = (mode & 64) ? (uint32_t)(i > 1) : 0u;  // 64: never start a fresh accumulation
#pragma unroll
      for (int k = 0; k < NB; ++k)
        umma_i8(d, da + (uint64_t)(2u * (k & 3)), db + (uint64_t)(2u * (k & 3)), idesc, k > 0 ? 1u : keep);
      if (!(mode & 128)) t_issue += clock64() - a0;
      if (mode & 32) {
        if (mode & 4) mbar_wait(b_ready, 0);
        if (mode & 8) mbar_wait(b_ready, 0);
        if (mode & 16) tc_fence_after();
      }
      if (mode & 1) umma_commit(b_commit);
      if (mode & 2) umma_commit(b_commit2);
      if (!(mode & 32)) {
        if (mode & 4) mbar_wait(b_ready, 0);
        if (mode & 8) mbar_wait(b_ready, 0);
        if (mode & 16) tc_fence_after();
      }
    }
    umma_commit(b_done);
    mbar_wait(b_done, 0);
    const long long t1 = clock64();
    // queue depth: from an idle pipe, how long does the thread take to issue the first j MMAs?
    long long q[6];
    {
      const long long s0 = clock64();
#pragma unroll
      for (int k = 0; k < 32; ++k) {
        umma_i8(tmem, da + (uint64_t)(2u * (k & 3)), db + (uint64_t)(2u * (k & 3)), idesc, 1u);
        if (k == 0) q[0] = clock64() - s0;
        if (k == 1) q[1] = clock64() - s0;
        if (k == 3) q[2] = clock64() - s0;
        if (k == 7) q[3] = clock64() - s0;
        if (k == 15) q[4] = clock64() - s0;
        if (k == 31) q[5] = clock64() - s0;
      }
      umma_commit(b_commit2);
    }
    if (blockIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t_issue;
      for (int j = 0; j < 6; ++j) out[2 + j] = q[j];
    }
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

int main(int argc, char** argv) {
  const int bn = argc > 1 ? atoi(argv[1]) : 64, iters = 2000;
  int nb = argc > 2 ? atoi(argv[2]) : 18;
  const int pix16 = argc > 3 ? atoi(argv[3]) : 0;
  long long* out;
  cudaMalloc(&out, 64);
  const size_t smem = (size_t)(160 + bn) * 128 + 1024;
  auto kern = nb == 2 ? sync_probe_kernel<2> : nb == 8 ? sync_probe_kernel<8> : nb == 9 ? sync_probe_kernel<9>
            : nb == 36 ? sync_probe_kernel<36> : sync_probe_kernel<18>;
  if (nb != 2 && nb != 8 && nb != 9 && nb != 36) nb = 18;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int modes[] = {0, 128, 192, 129, 131, 132, 140, 143, 159};
  printf("bn %d, %d MMAs per batch, A layout %s (back-to-back floor: see tools/mma_probe.py)\n", bn, nb, pix16 ? "pix16" : "SW128");
  for (int m : modes) {
    long long h[8] = {0};
    for (int rep = 0; rep < 2; ++rep) {
      kern<<<148, 128, smem>>>(iters, bn, m, out, pix16);
      if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("mode %d: %s\n", m, cudaGetErrorString(cudaGetLastError()));
        return 1;
      }
    }
    cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost);
    printf("mode %2d: %7.1f cycles per batch (%5.1f per MMA), %7.1f of them inside the issue loop\n", m,
           (double)h[0] / iters, (double)h[0] / iters / nb, (double)h[1] / iters);
    if (m == 0)
      printf("         issue of the first 1/2/4/8/16/32 MMAs into an idle pipe: %lld %lld %lld %lld %lld %lld cycles\n", h[2], h[3],
             h[4], h[5], h[6], h[7]);
  }
  return 0;
}
