// KernelContext / CUDAContext (reference: lite/core/context.h:441-448,
// lite/backends/cuda/context.h:34-167).  The int8 kernels only ever ENQUEUE on
// exec_stream(); io_copy uses io_stream().  One context per kernel instance.
#pragma once
#include <memory>

#include "lite/backends/cuda/target_wrapper.h"

namespace paddle {
namespace lite {

class CUDAContext {
 public:
  void Init(int dev_id, int exec_stream_id = 0, int io_stream_id = 0) {
    (void)exec_stream_id;
    (void)io_stream_id;
    device_id_ = dev_id;
  }
  void SetExecStream(cudaStream_t s) { exec_stream_ = s; }
  void SetIoStream(cudaStream_t s) { io_stream_ = s; }
  const cudaStream_t& exec_stream() const { return exec_stream_; }
  const cudaStream_t& io_stream() const { return io_stream_; }
  int device_id() const { return device_id_; }
  void SetNeedSync(bool s) { need_sync_ = s; }
  bool need_sync() const { return need_sync_; }
  void Sync() {
    if (need_sync_) TargetWrapperCuda::StreamSync(exec_stream_);
  }
  std::string name() const { return "CUDAContext"; }

 private:
  int device_id_{0};
  cudaStream_t exec_stream_{nullptr};
  cudaStream_t io_stream_{nullptr};
  bool need_sync_{false};
};

class KernelContext {
 public:
  template <typename ContextT>
  ContextT& As() {
    if (!cuda_) cuda_.reset(new CUDAContext);
    return *cuda_;
  }

 private:
  std::unique_ptr<CUDAContext> cuda_;
};

}  // namespace lite
}  // namespace paddle
