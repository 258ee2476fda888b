// conv2d / depthwise_conv2d kCUDA kInt8 kNHWC kernels.
// Replaces lite/kernels/cuda/conv_compute.{h,cc}:55-106,266-278 (ConvComputeInt8 over
// cuDNN) and adds the depthwise registration the reference lacks.  Input NHWC int8,
// filter OHWI int8 ([K,R,S,C/groups]), bias fp32, output NHWC int8 or fp32.
#pragma once
#include "lite/kernels/cuda/int8_common.h"

namespace paddle {
namespace lite {
namespace kernels {
namespace cuda {

template <PrecisionType Ptype_out>
class ConvComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::ConvParam;
  void PrepareForRun() override;
  void Run() override;
  virtual ~ConvComputeInt8() = default;

 private:
  enum Path { kDense, kDepthwise, kThinIm2col, kThinDirect, kThinS2d };
  Path path_{kDense};
  FoldedEpilogue epi_;
  Tensor packed_filter_;  // dense: tcgen05 K-blocked packing; thin: [K, ldA] or the direct-kernel packing
  Tensor x_padded_;       // input copied to a 16-byte channel pitch when C % 16 != 0
  Tensor im2col_;         // thin first layers (C <= 4)
  Tensor s2d_in_;         // stride-2 thin stems: the input as space-to-depth pixels (zero borders, written once)
  size_t s2d_bytes_{0};
  int ldA_{0};
};

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle
