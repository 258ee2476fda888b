// TargetWrapper<kCUDA>: device memory and copies (reference:
// lite/backends/cuda/target_wrapper.{h,cc}).  cudaMalloc / cudaMemcpy based, same
// entry points the reference's Tensor / io_copy kernels use.
#pragma once
#include <cuda_runtime.h>

#include "lite/core/types.h"
#include "lite/utils/logging.h"

namespace paddle {
namespace lite {

enum class IoDirection { HtoH = 0, HtoD, DtoH, DtoD };

template <TargetType T>
class TargetWrapper;

template <>
class TargetWrapper<TARGET(kHost)> {
 public:
  static void* Malloc(size_t size) { return size ? ::operator new(size) : nullptr; }
  static void Free(void* p) { ::operator delete(p); }
};

template <>
class TargetWrapper<TARGET(kCUDA)> {
 public:
  using stream_t = cudaStream_t;
  using event_t = cudaEvent_t;
  static size_t num_devices() {
    int n = 0;
    CUDA_CALL(cudaGetDeviceCount(&n));
    return static_cast<size_t>(n);
  }
  static int GetCurDevice() {
    int d = 0;
    CUDA_CALL(cudaGetDevice(&d));
    return d;
  }
  static void SetDevice(int id) { CUDA_CALL(cudaSetDevice(id)); }
  static void CreateStream(stream_t* s) { CUDA_CALL(cudaStreamCreate(s)); }
  static void DestroyStream(const stream_t& s) { CUDA_CALL(cudaStreamDestroy(s)); }
  static void StreamSync(const stream_t& s) { CUDA_CALL(cudaStreamSynchronize(s)); }
  static void DeviceSync() { CUDA_CALL(cudaDeviceSynchronize()); }
  static void* Malloc(size_t size) {
    void* p = nullptr;
    if (size) CUDA_CALL(cudaMalloc(&p, size));
    return p;
  }
  static void Free(void* p) {
    if (p) cudaFree(p);
  }
  static cudaMemcpyKind Kind(IoDirection dir) {
    switch (dir) {
      case IoDirection::HtoD: return cudaMemcpyHostToDevice;
      case IoDirection::DtoH: return cudaMemcpyDeviceToHost;
      case IoDirection::DtoD: return cudaMemcpyDeviceToDevice;
      default: return cudaMemcpyHostToHost;
    }
  }
  static void MemcpySync(void* dst, const void* src, size_t size, IoDirection dir) {
    if (size) CUDA_CALL(cudaMemcpy(dst, src, size, Kind(dir)));
  }
  static void MemcpyAsync(void* dst, const void* src, size_t size, IoDirection dir, const stream_t& stream) {
    if (size) CUDA_CALL(cudaMemcpyAsync(dst, src, size, Kind(dir), stream));
  }
  static void MemsetAsync(void* p, int v, size_t size, const stream_t& stream) {
    if (size) CUDA_CALL(cudaMemsetAsync(p, v, size, stream));
  }
};

using TargetWrapperCuda = TargetWrapper<TARGET(kCUDA)>;

}  // namespace lite
}  // namespace paddle
