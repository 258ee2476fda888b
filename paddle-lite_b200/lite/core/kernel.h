// KernelBase / KernelLite: the plugin surface (reference: lite/core/kernel.h:42-242).
// Launch() = first call PrepareForRun(), every call ReInitWhenNeeded() + Run().
#pragma once
#include <memory>
#include <string>
#include <typeindex>

#include "lite/core/context.h"
#include "lite/core/types.h"
#include "lite/utils/logging.h"

namespace paddle {
namespace lite {

// type-erased parameter slot (the reference uses lite::Any, lite/utils/any.h)
class ParamSlot {
 public:
  template <typename T>
  void set(const T& v) {
    ptr_ = std::make_shared<T>(v);
    type_ = std::type_index(typeid(T));
  }
  template <typename T>
  T* get_mutable() const {
    CHECK(ptr_ && type_ == std::type_index(typeid(T))) << "param type mismatch";
    return static_cast<T*>(ptr_.get());
  }
  bool valid() const { return static_cast<bool>(ptr_); }

 private:
  std::shared_ptr<void> ptr_;
  std::type_index type_{typeid(void)};
};

class KernelBase {
 public:
  virtual void PrepareForRun() {}
  virtual void ReInitWhenNeeded() {}
  virtual void Run() = 0;

  void Launch() {
    if (is_first_epoch_) {
      PrepareForRun();
      is_first_epoch_ = false;
    }
    ReInitWhenNeeded();
    Run();
  }
  void SetContext(std::unique_ptr<KernelContext>&& ctx) { ctx_ = std::move(ctx); }
  template <typename T>
  void SetParam(T param) {
    param_.set(param);
  }
  template <typename P>
  P& Param() const {
    return *param_.get_mutable<P>();
  }
  void set_op_type(const std::string& t) { op_type_ = t; }
  const std::string& op_type() const { return op_type_; }
  void set_alias(const std::string& a) { alias_ = a; }
  const std::string& alias() const { return alias_; }
  virtual Place place() const = 0;
  virtual TargetType target() const = 0;
  virtual PrecisionType precision() const = 0;
  virtual DataLayoutType layout() const = 0;
  virtual std::string name() const = 0;
  std::string key_with_alias() const { return op_type_ + "/" + alias_; }
  // "op/alias/target/precision/layout" (KernelBase::SerializeKernelType)
  std::string SerializedKernelType() const { return key_with_alias() + "/" + place().DebugString(); }
  KernelContext* mutable_context() { return ctx_.get(); }
  virtual ~KernelBase() = default;

 protected:
  std::unique_ptr<KernelContext> ctx_{nullptr};
  mutable ParamSlot param_;
  std::string op_type_{};
  std::string alias_{};
  bool is_first_epoch_{true};
};

template <TargetType Target, PrecisionType Precision, DataLayoutType DataLayout = DataLayoutType::kNCHW>
class KernelLite : public KernelBase {
 public:
  void Run() override { CHECK(false) << "Not Implemented"; }
  TargetType target() const override { return Target; }
  PrecisionType precision() const override { return Precision; }
  DataLayoutType layout() const override { return DataLayout; }
  Place place() const override { return Place{Target, Precision, DataLayout}; }
  std::string name() const override { return op_type() + ":" + place().DebugString(); }
  void Touch() {}
};

}  // namespace lite
}  // namespace paddle
