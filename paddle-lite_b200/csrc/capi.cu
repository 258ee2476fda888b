// capi.cu -- host-only pieces of the C-ABI (include/pl_int8.h): launch counter,
// output-size and scale-folding helpers.
//   fold rules: lite/kernels/arm/conv_gemmlike.cc:207-264,
//               lite/kernels/arm/fc_compute.cc:117-164,
//               lite/backends/cuda/math/cudnn_conv.cc:349-372 (which scales the
//               SHARED bias tensor in place; this layer writes private copies).
//   out size:   lite/operators/conv_op.cc:42-53
#include "common.cuh"

#include <cstdlib>

namespace pl {
std::atomic<unsigned long long> g_launches{0};

bool pdl_enabled() {
  static const bool on = [] {
    const char* v = getenv("PL_PDL");
    return !(v && v[0] == '0');
  }();
  return on;
}
}  // namespace pl

extern "C" int64_t pl_launch_count(void) { return (int64_t)pl::g_launches.load(std::memory_order_relaxed); }

extern "C" const char* pl_version(void) { return "paddle-lite_b200 int8 hot path 0.1 (sm_100a)"; }

extern "C" int pl_conv_out_size(int in, int k, int dil, int pad0, int pad1, int stride) {
  const int dkernel = dil * (k - 1) + 1;
  return (in + pad0 + pad1 - dkernel) / stride + 1;
}

extern "C" int pl_fold_scales(const float* w_scale, int n_w_scale, int n_out, float in_scale,
                              float out_scale, int out_dtype, const float* bias, float* scale_folded,
                              float* bias_folded, int act, float act_arg, float* act_arg_folded) {
  if (!w_scale || !scale_folded || n_out <= 0) return PL_ERR_ARG;
  if (n_w_scale != 1 && n_w_scale != n_out) return PL_ERR_ARG;
  for (int i = 0; i < n_out; ++i) {
    const float ws = w_scale[n_w_scale == 1 ? 0 : i];
    if (out_dtype == PL_OUT_INT8) {
      scale_folded[i] = ws * in_scale / out_scale;
      if (bias_folded) bias_folded[i] = bias ? bias[i] / out_scale : 0.f;
    } else {
      scale_folded[i] = ws * in_scale;
      if (bias_folded) bias_folded[i] = bias ? bias[i] : 0.f;
    }
  }
  if (act_arg_folded) {
    *act_arg_folded = act_arg;
    if (act == PL_ACT_RELU6 && out_dtype == PL_OUT_INT8) *act_arg_folded = act_arg / out_scale;
  }
  return PL_OK;
}
