// TMEM read bandwidth: W warps (one or more per lane quarter) read 16-column slices with tcgen05.ld.32x32b.x16
// back to back; prints cycles per load instruction and bytes per cycle per SM.  Optionally one thread keeps the tensor
// pipe busy with 128 x 64 x 32 MMAs meanwhile (does the MMA rate drop while the epilogue reads TMEM?).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I paddle-lite_b200/csrc -I include tools/ldtm_probe.cu -o tools/bin/ldtm_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "common.cuh"

using namespace pl;

__global__ void __launch_bounds__(640, 1) ldtm_probe_kernel(int iters, int read_warps, int with_mma, int bn, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ int stop;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (uint32_t i = threadIdx.x; i < (160u + 256u) * 128u / 4u; i += blockDim.x)
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x01010101u));
  fence_proxy_async();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
    stop = 0;
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  long long t_read = 0, t_mma = 0;
  uint32_t sink = 0;
  if (warp >= 4 && warp < 4 + read_warps) {
    const int quarter = warp & 3;
    const uint32_t addr = tmem + ((uint32_t)(quarter * 32) << 16);
    uint32_t r[16];
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      tmem_ld_32x32b_x16(addr + (uint32_t)((i * 16 + (warp >> 2) * 64) & 255), r);
      tmem_ld_wait();
      sink ^= r[0] ^ r[15];
    }
    t_read = clock64() - t0;
  } else if (warp == 0 && with_mma && elect_one()) {
    const uint32_t idesc = make_idesc_i8((uint32_t)bn);
    const uint64_t hi = make_smem_desc_base(128u);
    const uint64_t da = hi | (uint64_t)((base >> 4) & 0x3fff), db = hi | (uint64_t)(((base + 160 * 128) >> 4) & 0x3fff);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tmem + 256u + (uint32_t)((i & 1) * bn);
#pragma unroll
      for (int k = 0; k < 8; ++k) umma_i8(d, da + (uint64_t)(2u * (k & 3)), db + (uint64_t)(2u * (k & 3)), idesc, 1u);
    }
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    t_mma = clock64() - t0;
  }
  if (blockIdx.x == 0) {
    if (warp == 4 && lane == 0) out[0] = t_read;
    if (warp == 0 && with_mma && t_mma) out[1] = t_mma;
    if (sink == 0x12345678u) out[2] = sink;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

int main() {
  long long* out;
  cudaMalloc(&out, 64);
  const size_t smem = (size_t)(160 + 256) * 128 + 1024;
  cudaFuncSetAttribute(ldtm_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int iters = 4000;
  for (int with_mma = 0; with_mma <= 1; ++with_mma)
    for (int bn : {64, 128})
      for (int rw : {0, 1, 4, 8, 16}) {
        if (!with_mma && (rw == 0 || bn != 64)) continue;
        long long h[3] = {0, 0, 0};
        cudaMemset(out, 0, 64);
        for (int rep = 0; rep < 2; ++rep) {
          ldtm_probe_kernel<<<148, 640, smem>>>(iters, rw, with_mma, bn, out);
          if (cudaDeviceSynchronize() != cudaSuccess) {
            printf("error: %s\n", cudaGetErrorString(cudaGetLastError()));
            return 1;
          }
        }
        cudaMemcpy(h, out, 24, cudaMemcpyDeviceToHost);
        printf("read warps %2d, MMAs %s", rw, with_mma ? "on " : "off");
        if (with_mma) printf(" (N=%3d): %6.1f cycles per MMA", bn, (double)h[1] / (iters * 8.0));
        if (rw) printf("; %6.1f cycles per load and warp, %6.1f B/cycle/SM", (double)h[0] / iters, rw * 2048.0 * iters / (double)h[0]);
        printf("\n");
      }
  return 0;
}
