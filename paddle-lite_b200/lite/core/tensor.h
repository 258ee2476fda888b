// DDim / Tensor: the subset of lite/core/tensor.h the int8 kernels touch.
// Memory is owned by a lazily (re)allocated Buffer on the tensor's target
// (Buffer::ResetLazy, lite/core/memory.h): mutable_data<T>(TARGET(kCUDA)) only
// reallocates when the tensor grows.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <numeric>
#include <vector>

#include "lite/backends/cuda/target_wrapper.h"
#include "lite/core/types.h"

namespace paddle {
namespace lite {

class DDim {
 public:
  using value_type = int64_t;
  DDim() = default;
  explicit DDim(const std::vector<value_type>& d) : d_(d) {}
  DDim(std::initializer_list<value_type> d) : d_(d) {}
  size_t size() const { return d_.size(); }
  value_type operator[](size_t i) const { return d_.at(i); }
  value_type& operator[](size_t i) { return d_.at(i); }
  value_type production() const {
    value_type p = 1;
    for (auto v : d_) p *= v;
    return p;
  }
  value_type count(int s, int e) const {
    value_type p = 1;
    for (int i = s; i < e; ++i) p *= d_.at(i);
    return p;
  }
  DDim Slice(int s, int e) const { return DDim(std::vector<value_type>(d_.begin() + s, d_.begin() + e)); }
  DDim Flatten2D(int col) const { return DDim({count(0, col), count(col, static_cast<int>(size()))}); }
  const std::vector<value_type>& Vectorize() const { return d_; }
  bool operator==(const DDim& o) const { return d_ == o.d_; }
  bool operator!=(const DDim& o) const { return d_ != o.d_; }

 private:
  std::vector<value_type> d_;
};

class Buffer {
 public:
  ~Buffer() { Free(); }
  void ResetLazy(TargetType target, size_t size) {
    if (target != target_ || size > space_) {
      Free();
      data_ = target == TARGET(kCUDA) ? TargetWrapperCuda::Malloc(size) : TargetWrapper<TARGET(kHost)>::Malloc(size);
      target_ = target;
      space_ = size;
    }
  }
  void* data() const { return data_; }
  TargetType target() const { return target_; }
  size_t space() const { return space_; }
  // caller-owned memory (lite_api::Tensor::ShareExternalMemory, lite/api/paddle_api.h:62-65): never freed here;
  // the caller keeps it valid for as long as the tensor is used
  void ShareExternal(void* data, size_t size, TargetType target) {
    Free();
    data_ = data;
    space_ = size;
    target_ = target;
    external_ = true;
  }

 private:
  void Free() {
    if (data_ && !external_) {
      if (target_ == TARGET(kCUDA))
        TargetWrapperCuda::Free(data_);
      else
        TargetWrapper<TARGET(kHost)>::Free(data_);
    }
    data_ = nullptr;
    space_ = 0;
    external_ = false;
  }
  bool external_{false};
  void* data_{nullptr};
  size_t space_{0};
  TargetType target_{TARGET(kHost)};
};

class Tensor {
 public:
  Tensor() : buffer_(std::make_shared<Buffer>()) {}
  void Resize(const DDim& d) { dims_ = d; }
  void Resize(const std::vector<int64_t>& d) { dims_ = DDim(d); }
  const DDim& dims() const { return dims_; }
  int64_t numel() const { return dims_.production(); }
  TargetType target() const { return target_; }
  PrecisionType precision() const { return precision_; }
  void set_precision(PrecisionType p) { precision_ = p; }
  size_t memory_size() const { return memory_size_; }

  template <typename T>
  T* mutable_data(TargetType target = TARGET(kHost)) {
    precision_ = lite_api::PrecisionTypeTrait<T>::Type();
    target_ = target;
    memory_size_ = static_cast<size_t>(numel()) * sizeof(T);
    buffer_->ResetLazy(target, memory_size_);
    return static_cast<T*>(buffer_->data());
  }
  template <typename T>
  const T* data() const {
    return static_cast<const T*>(buffer_->data());
  }
  const void* raw_data() const { return buffer_->data(); }

  // host -> this tensor on `Target` (lite/core/tensor.h:106-113: Assign<DType, DimT, Target>)
  template <typename T, typename DimT, TargetType Target>
  void Assign(const T* src, const DimT& dims) {
    Resize(dims);
    T* dst = mutable_data<T>(Target);
    if (Target == TARGET(kCUDA))
      TargetWrapperCuda::MemcpySync(dst, src, memory_size_, IoDirection::HtoD);
    else
      std::memcpy(dst, src, memory_size_);
  }
  void ShareExternalMemory(void* data, size_t memory_size, TargetType target) {
    buffer_->ShareExternal(data, memory_size, target);
    target_ = target;
    memory_size_ = memory_size;
  }
  void ShareDataWith(const Tensor& o) {
    buffer_ = o.buffer_;
    dims_ = o.dims_;
    target_ = o.target_;
    precision_ = o.precision_;
    memory_size_ = o.memory_size_;
  }

 private:
  DDim dims_;
  std::shared_ptr<Buffer> buffer_;
  TargetType target_{TARGET(kHost)};
  PrecisionType precision_{PRECISION(kUnk)};
  size_t memory_size_{0};
};

}  // namespace lite
}  // namespace paddle
