// Helpers shared by the kCUDA/kInt8 kernel classes: private device buffers for the
// folded scale/bias vectors (the reference scales the SHARED bias tensor in place,
// lite/backends/cuda/math/cudnn_conv.cc:361-364 -- not idempotent; these kernels keep
// private copies like lite/kernels/arm/conv_gemmlike.cc:250-258), activation
// decoding, and the C-ABI include (the "lite/backends/cuda/math" int8 entry points).
#pragma once
#include <cstring>
#include <vector>

#include "include/pl_int8.h"
#include "lite/core/kernel.h"
#include "lite/core/tensor.h"
#include "lite/operators/op_params.h"

// C-ABI status codes of include/pl_int8.h are CHECKed like CUDA_CALL (lite/backends/cuda/cuda_utils.h)
#ifndef PL_CALL
#define PL_CALL(expr)                                                          \
  do {                                                                         \
    int rc__ = (expr);                                                         \
    CHECK_EQ(rc__, 0) << "pl_int8: " #expr " failed with status " << rc__;     \
  } while (0)
#endif

namespace paddle {
namespace lite {
namespace kernels {
namespace cuda {

inline void UploadFloats(Tensor* t, const std::vector<float>& v) {
  t->Assign<float, lite::DDim, TARGET(kCUDA)>(v.data(), DDim(std::vector<int64_t>{static_cast<int64_t>(v.size())}));
}

// payload of a kCUDA / host tensor -> host memory
inline void DownloadBytes(const Tensor* t, void* dst, size_t bytes) {
  if (t->target() == TARGET(kCUDA))
    TargetWrapperCuda::MemcpySync(dst, t->raw_data(), bytes, IoDirection::DtoH);
  else
    std::memcpy(dst, t->raw_data(), bytes);
}

inline std::vector<float> DownloadFloats(const Tensor* t) {
  std::vector<float> v(static_cast<size_t>(t->numel()));
  if (t->target() == TARGET(kCUDA))
    TargetWrapperCuda::MemcpySync(v.data(), t->raw_data(), v.size() * sizeof(float), IoDirection::DtoH);
  else
    std::memcpy(v.data(), t->raw_data(), v.size() * sizeof(float));
  return v;
}

inline int64_t RoundUp(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

struct FoldedEpilogue {
  Tensor scale, bias;
  int act{PL_ACT_NONE};
  float act_arg{0.f};

  // out_int8: fold /output_scale into scale, bias and the relu6 clip
  void Fold(const std::vector<float>& w_scale, int n_out, float in_scale, float out_scale, bool out_int8,
            const Tensor* bias_t, int act_, float act_arg_, float extra_mul = 1.f) {
    CHECK(w_scale.size() == 1 || static_cast<int>(w_scale.size()) == n_out)
        << "weight_scale must have 1 or " << n_out << " entries, got " << w_scale.size();
    std::vector<float> b;
    if (bias_t) {
      CHECK_EQ(bias_t->numel(), n_out);
      b = DownloadFloats(bias_t);
    }
    std::vector<float> ws(w_scale);
    if (extra_mul != 1.f)
      for (auto& s : ws) s *= extra_mul;
    std::vector<float> sf(n_out), bf(n_out);
    float arg = 0.f;
    PL_CALL(pl_fold_scales(ws.data(), static_cast<int>(ws.size()), n_out, in_scale, out_scale,
                           out_int8 ? PL_OUT_INT8 : PL_OUT_FP32, bias_t ? b.data() : nullptr, sf.data(), bf.data(),
                           act_, act_arg_, &arg));
    UploadFloats(&scale, sf);
    UploadFloats(&bias, bf);
    act = act_;
    act_arg = arg;
  }

  pl_epilogue Make(bool out_int8) const {
    pl_epilogue e{};
    e.scale = scale.data<float>();
    e.bias = bias.data<float>();
    e.act = act;
    e.act_arg = act_arg;
    e.out_dtype = out_int8 ? PL_OUT_INT8 : PL_OUT_FP32;
    e.round_mode = PL_ROUND_ARM;
    return e;
  }
};

inline void DecodeActivation(const operators::ActivationParam& a, int* act, float* arg) {
  *act = PL_ACT_NONE;
  *arg = 0.f;
  if (!a.has_active) return;
  switch (a.active_type) {
    case lite_api::ActivationType::kRelu: *act = PL_ACT_RELU; break;
    case lite_api::ActivationType::kRelu6: *act = PL_ACT_RELU6; *arg = a.Relu_clipped_coef; break;
    case lite_api::ActivationType::kLeakyRelu: *act = PL_ACT_LEAKY; *arg = a.Leaky_relu_alpha; break;
    default: CHECK(false) << "unsupported fused activation " << static_cast<int>(a.active_type);
  }
}

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle
