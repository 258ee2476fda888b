// Declared argument types of a kernel (reference: lite/core/type_system.h:299-336,
// ParamTypeRegistry).  static_kernel_pick selects between the int8_out / fp32_out
// registrations of an op by the DECLARED precision of "Output"/"Out", not by the
// alias string (lite/core/mir/static_kernel_pick_pass.cc:144-166).
#pragma once
#include <map>
#include <string>

#include "lite/core/types.h"

namespace paddle {
namespace lite {

struct Type {
  TargetType target{TARGET(kUnk)};
  PrecisionType precision{PRECISION(kUnk)};
  DataLayoutType layout{DATALAYOUT(kUnk)};
  Type() = default;
  Type(TargetType t, PrecisionType p, DataLayoutType l) : target(t), precision(p), layout(l) {}
  // lite/core/type_system.h:107-118 (returns a const Type* there; the kernel sources only ever write
  // `{LiteType::GetTensorTy(...)}` inside BindInput / BindOutput, which both forms accept)
  static Type GetTensorTy(TargetType t, PrecisionType p = PRECISION(kFloat), DataLayoutType l = DATALAYOUT(kNCHW),
                          int device = 0) {
    (void)device;
    return Type(t, p, l);
  }
};

class ParamTypeRegistry {
 public:
  struct KernelDecl {
    Place place;
    std::map<std::string, Type> inputs, outputs;
  };
  // Fluent recorder returned by REGISTER_LITE_KERNEL(...)
  class NewInstanceT {
   public:
    NewInstanceT(const std::string& key, Place place) : key_(key) { Global().decls_[key_].place = place; }
    NewInstanceT& BindInput(const std::string& arg, std::initializer_list<Type> tys) {
      Global().decls_[key_].inputs[arg] = *tys.begin();
      return *this;
    }
    NewInstanceT& BindOutput(const std::string& arg, std::initializer_list<Type> tys) {
      Global().decls_[key_].outputs[arg] = *tys.begin();
      return *this;
    }
    bool Finalize() { return true; }

   private:
    std::string key_;
  };
  template <TargetType T, PrecisionType P, DataLayoutType L>
  static NewInstanceT NewInstance(const std::string& key) {
    return NewInstanceT(key + "/" + Place(T, P, L).DebugString(), Place(T, P, L));
  }
  static ParamTypeRegistry& Global() {
    static ParamTypeRegistry r;
    return r;
  }
  const KernelDecl* Find(const std::string& op_alias, const Place& place) const {
    auto it = decls_.find(op_alias + "/" + place.DebugString());
    return it == decls_.end() ? nullptr : &it->second;
  }
  const std::map<std::string, KernelDecl>& all() const { return decls_; }

 private:
  std::map<std::string, KernelDecl> decls_;
};

}  // namespace lite
}  // namespace paddle
