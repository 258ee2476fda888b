// StaticKernelPickPass for one statement, restated (reference: lite/core/mir/static_kernel_pick_pass.{h,cc},
// candidate creation lite/core/op_lite.cc:82-126).  What is reproduced, step by step:
//   1. candidates = every kernel registered for the op on the valid places EXPANDED by their kAny-precision /
//      kAny-layout variants, visited in std::set<Place> order (paddle_place.cc:33-39: target, precision, layout)
//      -- op_lite.cc:107-122;
//   2. KernelGrade (static_kernel_pick_pass.h:53-215): for valid place i, weight (n - i) / n times the sum of
//      kMax / 1 (target matches), kMax / 2 (precision matches; an int8 kernel scores nothing here unless the op
//      has enable_int8), kMax / 4 (layout matches), kMax = 255 (KernelPickFactor::value_type = unsigned char,
//      integer division); for an op WITHOUT enable_int8 the sum doubles when every known input precision is
//      compatible with the kernel's declaration (fp16 ~ float); the best place wins; the float result is
//      returned through size_t, i.e. truncated;
//   3. stable sort by score, descending (cc:30-33,83): ties keep the candidate order of step 1;
//   4. no enable_int8: the top candidate (cc:86-91).  enable_int8: the output is int8 iff EVERY consumer has
//      enable_int8; then the first consumer's input scale becomes this op's output scale (cc:93-125) and the
//      first candidate in score order whose DECLARED output precisions all equal int8 (else float) is taken
//      (cc:144-166).
// The FPGA special case (h:146-183) is not restated: no such kernels exist here.
#pragma once
#include <algorithm>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "lite/core/op_registry.h"

namespace paddle {
namespace lite {
namespace mir {

struct ConsumerInfo {
  bool enable_int8{false};
  float input_scale{1.f};
};

struct PlaceLess {  // lite/api/paddle_place.cc:33-39
  bool operator()(const Place& a, const Place& b) const {
    if (a.target != b.target) return a.target < b.target;
    if (a.precision != b.precision) return a.precision < b.precision;
    if (a.layout != b.layout) return a.layout < b.layout;
    return a.device < b.device;
  }
};

inline bool PrecTypeCompatible(PrecisionType p1, PrecisionType p2) {  // static_kernel_pick_pass.h:219-228
  if (p1 == p2) return true;
  const bool f1 = p1 == PRECISION(kFP16) || p1 == PRECISION(kFloat);
  const bool f2 = p2 == PRECISION(kFP16) || p2 == PRECISION(kFloat);
  return f1 && f2;
}

// A candidate as the grading rule sees it: the place it is registered on and its declared argument precisions
struct KernelView {
  Place place;
  const ParamTypeRegistry::KernelDecl* decl{nullptr};
};

// static_kernel_pick_pass.h:53-215.  in_types / out_types: precision of the statement's input / output
// variables where known, keyed by ARGUMENT name (the reference maps variable -> argument name first).
inline size_t KernelGrade(const KernelView& kernel, const std::vector<Place>& places, bool enable_int8, bool is_feed,
                          const std::map<std::string, PrecisionType>& in_types,
                          const std::map<std::string, PrecisionType>& out_types) {
  float final_score = -1.f;
  const int kMax = 255;  // std::numeric_limits<unsigned char>::max()
  const size_t place_size = places.size();
  for (size_t i = 0; i < place_size; ++i) {
    const Place& place = places[i];
    const float weight = static_cast<float>(place_size - i) / place_size;
    size_t score = 0;
    if (place.target == kernel.place.target || kernel.place.target == TARGET(kAny) || place.target == TARGET(kAny))
      score += kMax / 1;  // Factor::TargetFirst
    if (place.precision == kernel.place.precision || kernel.place.precision == PRECISION(kAny) ||
        place.precision == PRECISION(kAny)) {
      if (!(kernel.place.precision == PRECISION(kInt8) && !enable_int8)) score += kMax / 2;  // PrecisionFirst
    }
    if (place.layout == kernel.place.layout || kernel.place.layout == DATALAYOUT(kAny) ||
        place.layout == DATALAYOUT(kAny))
      score += kMax / 4;  // DataLayoutFirst
    if (!enable_int8) {
      bool type_match = true;
      if (kernel.decl) {
        const auto& known = is_feed ? out_types : in_types;
        const auto& declared = is_feed ? kernel.decl->outputs : kernel.decl->inputs;
        for (const auto& kv : known) {
          auto it = declared.find(kv.first);
          if (it == declared.end()) continue;
          const bool ok = is_feed ? kv.second == it->second.precision : PrecTypeCompatible(kv.second, it->second.precision);
          if (!ok) type_match = false;
        }
      }
      if (type_match) score *= 2;
    }
    if (weight * score > final_score) final_score = weight * score;
  }
  return static_cast<size_t>(final_score);
}

struct Scored {
  float score;
  std::unique_ptr<KernelBase> kernel;
};

// op_lite.cc:82-126 + static_kernel_pick_pass.cc:66-84: every candidate with its score, best first
inline std::vector<Scored> ScoreKernels(const std::string& op_type, const std::vector<Place>& valid_places,
                                        bool enable_int8, const std::map<std::string, PrecisionType>& in_types = {},
                                        const std::map<std::string, PrecisionType>& out_types = {}) {
  std::set<Place, PlaceLess> expanded(valid_places.begin(), valid_places.end());
  for (const auto& p : valid_places) {
    expanded.insert(Place(p.target, p.precision, DATALAYOUT(kAny)));
    expanded.insert(Place(p.target, PRECISION(kAny), p.layout));
    expanded.insert(Place(p.target, PRECISION(kAny), DATALAYOUT(kAny)));
  }
  std::vector<Scored> scored;
  for (const auto& place : expanded) {
    auto ks = KernelRegistry::Global().Create(op_type, place.target, place.precision, place.layout);
    for (auto& k : ks) {
      KernelView v;
      v.place = k->place();
      v.decl = ParamTypeRegistry::Global().Find(k->key_with_alias(), k->place());
      const float s = static_cast<float>(KernelGrade(v, valid_places, enable_int8, op_type == "feed", in_types, out_types));
      scored.push_back(Scored{s, std::move(k)});
    }
  }
  std::stable_sort(scored.begin(), scored.end(), [](const Scored& a, const Scored& b) { return a.score > b.score; });
  return scored;
}

inline std::unique_ptr<KernelBase> StaticKernelPick(const std::string& op_type, const std::vector<Place>& valid_places,
                                                    bool enable_int8, const std::vector<ConsumerInfo>& consumers,
                                                    float* output_scale = nullptr,
                                                    const std::map<std::string, PrecisionType>& in_types = {},
                                                    const std::map<std::string, PrecisionType>& out_types = {}) {
  auto scored = ScoreKernels(op_type, valid_places, enable_int8, in_types, out_types);
  if (scored.empty()) return nullptr;
  if (!enable_int8) {
    // an int8 kernel can still sit on top for a non-quantised op only through target + layout; the reference keeps
    // whatever is first (cc:86-91) -- but it never registers int8-only ops, so mirror its outcome: no fp32 kernel
    // registered here -> nothing to pick
    for (auto& c : scored)
      if (c.kernel->place().precision != PRECISION(kInt8)) return std::move(c.kernel);
    return nullptr;
  }
  bool out_int8 = !consumers.empty();
  for (const auto& c : consumers) out_int8 = out_int8 && c.enable_int8;
  if (out_int8 && output_scale) *output_scale = consumers.front().input_scale;
  const PrecisionType expect = out_int8 ? PRECISION(kInt8) : PRECISION(kFloat);
  for (auto& c : scored) {
    const auto* decl = ParamTypeRegistry::Global().Find(c.kernel->key_with_alias(), c.kernel->place());
    if (!decl || decl->outputs.empty()) continue;
    bool all_match = true;
    for (const auto& kv : decl->outputs)
      if (kv.second.precision != expect) all_match = false;
    if (all_match) return std::move(c.kernel);
  }
  return nullptr;
}

}  // namespace mir
}  // namespace lite
}  // namespace paddle
