// Place / enum vocabulary of the plugin boundary.  Values are the reference's
// (lite/api/paddle_place.h:45-114) so that serialized kernel keys stay compatible.
#pragma once
#include <cstdint>
#include <string>

namespace paddle {
namespace lite_api {

enum class TargetType : int { kUnk = 0, kHost = 1, kX86 = 2, kCUDA = 3, kARM = 4, kAny = 6 };
enum class PrecisionType : int {
  kUnk = 0, kFloat = 1, kInt8 = 2, kInt32 = 3, kAny = 4, kFP16 = 5, kBool = 6, kInt64 = 7, kInt16 = 8, kUInt8 = 9
};
enum class DataLayoutType : int { kUnk = 0, kNCHW = 1, kAny = 2, kNHWC = 3 };
enum class ActivationType : int { kIndentity = 0, kRelu = 1, kRelu6 = 2, kPRelu = 3, kLeakyRelu = 4 };

inline size_t PrecisionTypeLength(PrecisionType t) {
  switch (t) {
    case PrecisionType::kFloat: case PrecisionType::kInt32: return 4;
    case PrecisionType::kInt8: case PrecisionType::kUInt8: case PrecisionType::kBool: return 1;
    case PrecisionType::kFP16: case PrecisionType::kInt16: return 2;
    case PrecisionType::kInt64: return 8;
    default: return 0;
  }
}

inline const char* TargetToStr(TargetType t) {
  switch (t) {
    case TargetType::kHost: return "host";
    case TargetType::kX86: return "x86";
    case TargetType::kCUDA: return "cuda";
    case TargetType::kARM: return "arm";
    case TargetType::kAny: return "any";
    default: return "unk";
  }
}
inline const char* PrecisionToStr(PrecisionType p) {
  switch (p) {
    case PrecisionType::kFloat: return "float";
    case PrecisionType::kInt8: return "int8_t";
    case PrecisionType::kInt32: return "int32_t";
    case PrecisionType::kAny: return "any";
    case PrecisionType::kFP16: return "float16";
    default: return "unk";
  }
}
inline const char* DataLayoutToStr(DataLayoutType l) {
  switch (l) {
    case DataLayoutType::kNCHW: return "NCHW";
    case DataLayoutType::kNHWC: return "NHWC";
    case DataLayoutType::kAny: return "any";
    default: return "unk";
  }
}

// lite/api/paddle_place.h:147-175
template <typename T>
struct PrecisionTypeTrait {
  constexpr static PrecisionType Type() { return PrecisionType::kUnk; }
};
#define PL_DEFINE_PRECISION_TRAIT(cpp_type, precision_type)                         \
  template <>                                                                      \
  struct PrecisionTypeTrait<cpp_type> {                                            \
    constexpr static PrecisionType Type() { return PrecisionType::precision_type; } \
  }
PL_DEFINE_PRECISION_TRAIT(bool, kBool);
PL_DEFINE_PRECISION_TRAIT(float, kFloat);
PL_DEFINE_PRECISION_TRAIT(uint8_t, kUInt8);
PL_DEFINE_PRECISION_TRAIT(int8_t, kInt8);
PL_DEFINE_PRECISION_TRAIT(int16_t, kInt16);
PL_DEFINE_PRECISION_TRAIT(int, kInt32);
PL_DEFINE_PRECISION_TRAIT(int64_t, kInt64);
#undef PL_DEFINE_PRECISION_TRAIT

struct Place {
  TargetType target{TargetType::kUnk};
  PrecisionType precision{PrecisionType::kUnk};
  DataLayoutType layout{DataLayoutType::kUnk};
  int16_t device{0};
  Place() = default;
  Place(TargetType t, PrecisionType p = PrecisionType::kFloat, DataLayoutType l = DataLayoutType::kNCHW, int16_t d = 0)
      : target(t), precision(p), layout(l), device(d) {}
  bool operator==(const Place& o) const {
    return target == o.target && precision == o.precision && layout == o.layout && device == o.device;
  }
  std::string DebugString() const {
    return std::string(TargetToStr(target)) + "/" + PrecisionToStr(precision) + "/" + DataLayoutToStr(layout);
  }
};

}  // namespace lite_api

namespace lite {
using lite_api::TargetType;
using lite_api::PrecisionType;
using lite_api::DataLayoutType;
using lite_api::ActivationType;
using lite_api::Place;
}  // namespace lite
}  // namespace paddle

#define TARGET(item__) ::paddle::lite_api::TargetType::item__
#define PRECISION(item__) ::paddle::lite_api::PrecisionType::item__
#define DATALAYOUT(item__) ::paddle::lite_api::DataLayoutType::item__
