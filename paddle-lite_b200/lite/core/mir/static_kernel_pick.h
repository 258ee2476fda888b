// Miniature of StaticKernelPickPass for one op (reference:
// lite/core/mir/static_kernel_pick_pass.{h,cc}).  Rules reproduced:
//   * place order = preference: the first valid place that has a kernel wins
//     (static_kernel_pick_pass.h:67-78);
//   * int8 kernels are only eligible when the op has enable_int8 (h:93-98);
//   * for an enable_int8 op the output is int8 iff EVERY consumer also has enable_int8
//     (cc:93-106); the candidate whose DECLARED "Output"/"Out" precision matches is
//     picked -- int8 -> the int8_out registration, else fp32_out (cc:144-166);
//   * the consumer's input scale becomes this op's output_scale (cc:110-125) -- returned
//     to the caller through `output_scale`.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "lite/core/op_registry.h"

namespace paddle {
namespace lite {
namespace mir {

struct ConsumerInfo {
  bool enable_int8{false};
  float input_scale{1.f};
};

inline std::unique_ptr<KernelBase> StaticKernelPick(const std::string& op_type, const std::vector<Place>& valid_places,
                                                    bool enable_int8, const std::vector<ConsumerInfo>& consumers,
                                                    float* output_scale = nullptr) {
  bool out_int8 = enable_int8 && !consumers.empty();
  for (const auto& c : consumers) out_int8 = out_int8 && c.enable_int8;
  if (out_int8 && output_scale) *output_scale = consumers.front().input_scale;
  const PrecisionType want_out = out_int8 ? PRECISION(kInt8) : PRECISION(kFloat);
  for (const auto& place : valid_places) {
    if (place.precision == PRECISION(kInt8) && !enable_int8) continue;
    auto kernels = KernelRegistry::Global().Create(op_type, place.target, place.precision, place.layout);
    for (auto& k : kernels) {
      const auto* decl = ParamTypeRegistry::Global().Find(k->key_with_alias(), k->place());
      if (!decl) continue;
      auto it = decl->outputs.find("Output");
      if (it == decl->outputs.end()) it = decl->outputs.find("Out");
      if (it == decl->outputs.end()) continue;
      if (place.precision != PRECISION(kInt8) || it->second.precision == want_out) return std::move(k);
    }
  }
  return nullptr;
}

}  // namespace mir
}  // namespace lite
}  // namespace paddle
