// Kernel registration (reference: lite/core/op_registry.h:125-254).
// REGISTER_LITE_KERNEL(op, target, precision, layout, Class, alias)
//     .BindInput(...).BindOutput(...).Finalize();
// inserts a creator keyed by (op, target, precision, layout) into
// KernelRegistry::Global() at static-init time and records the declared
// argument types in ParamTypeRegistry; it also defines touch_<...>() so that a
// binary can force-link the registration (USE_LITE_KERNEL).
#pragma once
#include <functional>
#include <list>
#include <map>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#include "lite/core/kernel.h"
#include "lite/core/type_system.h"

using LiteType = paddle::lite::Type;  // lite/core/op_registry.h:31

namespace paddle {
namespace lite {

class KernelRegistry {
 public:
  using creator_t = std::function<std::unique_ptr<KernelBase>()>;
  using key_t = std::tuple<std::string, int, int, int>;
  static KernelRegistry& Global() {
    static KernelRegistry r;
    return r;
  }
  void Register(const std::string& op, TargetType t, PrecisionType p, DataLayoutType l, creator_t c) {
    creators_[key_t(op, (int)t, (int)p, (int)l)].push_back(std::move(c));
  }
  // every kernel registered for (op, place): one per alias
  std::list<std::unique_ptr<KernelBase>> Create(const std::string& op, TargetType t, PrecisionType p,
                                                DataLayoutType l) const {
    std::list<std::unique_ptr<KernelBase>> out;
    auto it = creators_.find(key_t(op, (int)t, (int)p, (int)l));
    if (it != creators_.end())
      for (auto& c : it->second) out.push_back(c());
    return out;
  }
  std::vector<std::string> DebugKeys() const {
    std::vector<std::string> ks;
    for (auto& kv : creators_)
      for (auto& c : kv.second) ks.push_back(c()->SerializedKernelType());
    return ks;
  }

 private:
  std::map<key_t, std::list<creator_t>> creators_;
};

class KernelRegistrar {
 public:
  KernelRegistrar(const std::string& op, TargetType t, PrecisionType p, DataLayoutType l,
                  KernelRegistry::creator_t c) {
    KernelRegistry::Global().Register(op, t, p, l, std::move(c));
  }
  void touch() {}
};

}  // namespace lite
}  // namespace paddle

#define LITE_UNUSED __attribute__((unused))
#define REGISTER_LITE_KERNEL(op_type__, target__, precision__, layout__, KernelClass, alias__)            \
  static paddle::lite::KernelRegistrar op_type__##target__##precision__##layout__##alias__##_kernel_registry( \
      #op_type__, TARGET(target__), PRECISION(precision__), DATALAYOUT(layout__), []() {                  \
        std::unique_ptr<paddle::lite::KernelBase> x(new KernelClass);                                     \
        x->set_op_type(#op_type__);                                                                       \
        x->set_alias(#alias__);                                                                           \
        return x;                                                                                         \
      });                                                                                                 \
  int touch_##op_type__##target__##precision__##layout__##alias__() {                                     \
    op_type__##target__##precision__##layout__##alias__##_kernel_registry.touch();                        \
    return 0;                                                                                             \
  }                                                                                                       \
  static auto op_type__##target__##precision__##layout__##alias__##param_register LITE_UNUSED =           \
      paddle::lite::ParamTypeRegistry::NewInstance<TARGET(target__), PRECISION(precision__),              \
                                                   DATALAYOUT(layout__)>(#op_type__ "/" #alias__)

#define USE_LITE_KERNEL(op_type__, target__, precision__, layout__, alias__)        \
  extern int touch_##op_type__##target__##precision__##layout__##alias__();         \
  int op_type__##target__##precision__##layout__##alias__##__use_lite_kernel LITE_UNUSED = \
      touch_##op_type__##target__##precision__##layout__##alias__();
