// calib / layout / io_copy / pool2d / elementwise_add kCUDA kernels of the int8 region.
//   calib      lite/kernels/cuda/calib_compute.cu:27-151 (regs :153-276)
//   layout     lite/kernels/cuda/layout_compute.cc:41-94 (regs to :222)
//   io_copy    lite/kernels/cuda/io_copy_compute.cc
//   pool2d     int8 NHWC: new (reference pools in fp32, lite/kernels/cuda/pool_compute.cu)
//   elementwise_add int8: new (lite/backends/cuda/math/elementwise.cu:127-165 is dead code)
// so that activations stay int8 + NHWC + on device between consecutive int8 ops.
#include "lite/core/op_registry.h"
#include "lite/kernels/cuda/int8_common.h"

namespace paddle {
namespace lite {
namespace kernels {
namespace cuda {

static pl_stream_t StreamOf(KernelContext* ctx) {
  return reinterpret_cast<pl_stream_t>(ctx->As<CUDAContext>().exec_stream());
}

class CalibComputeFp32ToInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::CalibParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    param.output->Resize(param.input->dims());
    PL_CALL(pl_quantize_f32_i8(param.input->data<float>(), param.output->mutable_data<int8_t>(TARGET(kCUDA)),
                               param.input->numel(), param.scale, PL_ROUND_ARM, StreamOf(ctx_.get())));
  }
};

class CalibComputeInt8ToFp32 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::CalibParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    param.output->Resize(param.input->dims());
    PL_CALL(pl_dequantize_i8_f32(param.input->data<int8_t>(), param.output->mutable_data<float>(TARGET(kCUDA)),
                                 param.input->numel(), param.scale, StreamOf(ctx_.get())));
  }
};

template <typename T, bool ToNHWC>
class LayoutCompute : public KernelLite<TARGET(kCUDA), lite_api::PrecisionTypeTrait<T>::Type(), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::LayoutParam;
  void Run() override {
    auto& param = this->template Param<param_t>();
    const auto& d = param.x->dims();
    CHECK_EQ(d.size(), 4u) << "layout expects a 4-D tensor";
    if (ToNHWC) {
      const int N = (int)d[0], C = (int)d[1], H = (int)d[2], W = (int)d[3];
      param.y->Resize(DDim({N, H, W, C}));
      PL_CALL(pl_nchw_to_nhwc(param.x->template data<T>(), param.y->template mutable_data<T>(TARGET(kCUDA)), N, C, H,
                              W, C, sizeof(T), StreamOf(this->ctx_.get())));
    } else {
      const int N = (int)d[0], H = (int)d[1], W = (int)d[2], C = (int)d[3];
      param.y->Resize(DDim({N, C, H, W}));
      PL_CALL(pl_nhwc_to_nchw(param.x->template data<T>(), param.y->template mutable_data<T>(TARGET(kCUDA)), N, C, H,
                              W, C, sizeof(T), StreamOf(this->ctx_.get())));
    }
  }
};

class IoCopyHostToCuda : public KernelLite<TARGET(kCUDA), PRECISION(kAny), DATALAYOUT(kAny)> {
 public:
  using param_t = operators::IoCopyParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    CHECK(param.x->target() == TARGET(kHost));
    param.y->Resize(param.x->dims());
    void* dst = Alloc(param.y, param.x->precision());
    TargetWrapperCuda::MemcpyAsync(dst, param.x->raw_data(), param.x->memory_size(), IoDirection::HtoD,
                                   ctx_->As<CUDAContext>().exec_stream());
  }
  static void* Alloc(Tensor* t, PrecisionType p) {
    switch (p) {
      case PRECISION(kInt8): return t->mutable_data<int8_t>(TARGET(kCUDA));
      case PRECISION(kInt32): return t->mutable_data<int32_t>(TARGET(kCUDA));
      default: return t->mutable_data<float>(TARGET(kCUDA));
    }
  }
};

class IoCopyCudaToHost : public KernelLite<TARGET(kCUDA), PRECISION(kAny), DATALAYOUT(kAny)> {
 public:
  using param_t = operators::IoCopyParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    CHECK(param.x->target() == TARGET(kCUDA));
    param.y->Resize(param.x->dims());
    void* dst;
    switch (param.x->precision()) {
      case PRECISION(kInt8): dst = param.y->mutable_data<int8_t>(TARGET(kHost)); break;
      case PRECISION(kInt32): dst = param.y->mutable_data<int32_t>(TARGET(kHost)); break;
      default: dst = param.y->mutable_data<float>(TARGET(kHost)); break;
    }
    // D2H is synchronous in the reference too (MemcpySync, io_copy_compute.cc:26-40)
    auto s = ctx_->As<CUDAContext>().exec_stream();
    TargetWrapperCuda::MemcpyAsync(dst, param.x->raw_data(), param.x->memory_size(), IoDirection::DtoH, s);
    TargetWrapperCuda::StreamSync(s);
  }
};

template <PrecisionType Ptype_out>
class PoolComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::PoolParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    const auto& d = param.x->dims();  // NHWC
    CHECK_EQ(d.size(), 4u);
    CHECK(param.pooling_type == "max" || param.pooling_type == "avg") << "pooling_type " << param.pooling_type;
    const int N = (int)d[0], H = (int)d[1], W = (int)d[2], C = (int)d[3];
    if (param.adaptive && !param.global_pooling) {
      // ksize is the output size (lite/kernels/cuda/pool_compute.cu:256-350, lite/operators/pool_op.cc)
      const int P = param.ksize.at(0), Q = param.ksize.at(1);
      param.output->Resize(DDim({N, P, Q, C}));
      const bool o8 = Ptype_out == PRECISION(kInt8);
      void* ya = o8 ? static_cast<void*>(param.output->template mutable_data<int8_t>(TARGET(kCUDA)))
                    : static_cast<void*>(param.output->template mutable_data<float>(TARGET(kCUDA)));
      PL_CALL(pl_adaptive_pool2d_i8_nhwc(param.x->template data<int8_t>(), C, ya, C, N, H, W, C, P, Q,
                                         param.pooling_type == "max" ? PL_POOL_MAX : PL_POOL_AVG,
                                         o8 ? param.input_scale / param.output_scale : param.input_scale,
                                         o8 ? PL_OUT_INT8 : PL_OUT_FP32, PL_ROUND_ARM, StreamOf(this->ctx_.get())));
      return;
    }
    int kh = param.ksize.at(0), kw = param.ksize.at(1);
    int pt = 0, pb = 0, pl = 0, pr = 0;
    if (param.paddings && param.paddings->size() == 4) {
      pt = (*param.paddings)[0]; pb = (*param.paddings)[1]; pl = (*param.paddings)[2]; pr = (*param.paddings)[3];
    }
    if (param.global_pooling) { kh = H; kw = W; pt = pb = pl = pr = 0; }
    const int sh = param.strides[0], sw = param.strides[1];
    int P, Q;
    if (param.ceil_mode) {
      P = (H + pt + pb - kh + sh - 1) / sh + 1;
      Q = (W + pl + pr - kw + sw - 1) / sw + 1;
    } else {
      P = (H + pt + pb - kh) / sh + 1;
      Q = (W + pl + pr - kw) / sw + 1;
    }
    param.output->Resize(DDim({N, P, Q, C}));
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    void* y = out_int8 ? static_cast<void*>(param.output->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : static_cast<void*>(param.output->template mutable_data<float>(TARGET(kCUDA)));
    const float rescale = out_int8 ? param.input_scale / param.output_scale : param.input_scale;
    PL_CALL(pl_pool2d_i8_nhwc(param.x->template data<int8_t>(), C, y, C, N, H, W, C, kh, kw, sh, sw, pt, pl, P, Q,
                              param.pooling_type == "max" ? PL_POOL_MAX : PL_POOL_AVG, param.exclusive ? 1 : 0,
                              rescale, out_int8 ? PL_OUT_INT8 : PL_OUT_FP32, PL_ROUND_ARM, StreamOf(ctx_.get())));
  }
};

// elementwise_add (ElementwiseParam) and fusion_elementwise_add_activation
// (FusionElementwiseActivationParam: + act_type), lite/operators/op_params.h:643-680
template <PrecisionType Ptype_out, typename ParamT>
class ElementwiseAddInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = ParamT;
  static int ActOf(const operators::ElementwiseParam&) { return PL_ACT_NONE; }
  static int ActOf(const operators::FusionElementwiseActivationParam& p) {
    CHECK(p.act_type == "" || p.act_type == "relu") << "int8 fused elementwise_add: relu only";
    return p.act_type == "relu" ? PL_ACT_RELU : PL_ACT_NONE;
  }
  void Run() override {
    auto& param = this->template Param<param_t>();
    CHECK(param.X->dims() == param.Y->dims()) << "int8 elementwise_add needs equal shapes (residual add)";
    param.Out->Resize(param.X->dims());
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    void* y = out_int8 ? static_cast<void*>(param.Out->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : static_cast<void*>(param.Out->template mutable_data<float>(TARGET(kCUDA)));
    const float sa = out_int8 ? param.x_input_scale / param.output_scale : param.x_input_scale;
    const float sb = out_int8 ? param.y_input_scale / param.output_scale : param.y_input_scale;
    PL_CALL(pl_add_i8(param.X->template data<int8_t>(), param.Y->template data<int8_t>(), y, param.X->numel(), sa, sb,
                      ActOf(param), 0.f, out_int8 ? PL_OUT_INT8 : PL_OUT_FP32, PL_ROUND_ARM,
                      StreamOf(this->ctx_.get())));
  }
};


// ---- channel movement between int8 NHWC ops (no calib / layout round trip inside a network) ----
// The op descs keep the model's NCHW axis numbering: channels are axis 1 (or -3); the tensors
// themselves are NHWC, so the reference's per-image memcpy loops (lite/backends/arm/math/concat.h:29-60,
// split.cc:54-83, shuffle_channel.cc:24-76) become per-pixel byte moves in the library.
inline bool IsChannelAxis(int axis) { return axis == 1 || axis == -3; }

class ConcatComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::ConcatParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    CHECK(IsChannelAxis(param.axis)) << "int8 NHWC concat: channel axis only, got axis " << param.axis;
    CHECK(!param.x.empty() && param.x.size() <= 8u) << "int8 NHWC concat takes 1..8 inputs";
    const auto& d0 = param.x[0]->dims();
    CHECK_EQ(d0.size(), 4u);
    std::vector<const int8_t*> ptrs;
    std::vector<int> cs;
    int64_t ctot = 0;
    for (auto* t : param.x) {
      const auto& d = t->dims();
      CHECK(d.size() == 4u && d[0] == d0[0] && d[1] == d0[1] && d[2] == d0[2]) << "concat inputs differ outside the channel axis";
      ptrs.push_back(t->template data<int8_t>());
      cs.push_back(static_cast<int>(d[3]));
      ctot += d[3];
    }
    param.output->Resize(DDim({d0[0], d0[1], d0[2], ctot}));
    auto* y = param.output->template mutable_data<int8_t>(TARGET(kCUDA));
    PL_CALL(pl_concat_i8_nhwc(static_cast<int>(ptrs.size()), ptrs.data(), cs.data(), cs.data(), y, static_cast<int>(ctot),
                              d0[0] * d0[1] * d0[2], StreamOf(ctx_.get())));
  }
};

class SplitComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::SplitParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    CHECK(IsChannelAxis(param.axis)) << "int8 NHWC split: channel axis only, got axis " << param.axis;
    const auto& d = param.x->dims();
    CHECK_EQ(d.size(), 4u);
    const int C = static_cast<int>(d[3]);
    std::vector<int> cs = param.sections;
    if (cs.empty()) {  // `num` equal parts (lite/operators/split_op.cc InferShape)
      CHECK(param.num > 0 && C % param.num == 0) << "split: num " << param.num << " does not divide C " << C;
      cs.assign(param.num, C / param.num);
    }
    CHECK(cs.size() == param.output.size() && cs.size() <= 8u) << "split: sections vs outputs";
    int sum = 0;
    std::vector<int8_t*> ptrs;
    for (size_t k = 0; k < cs.size(); ++k) {
      sum += cs[k];
      param.output[k]->Resize(DDim({d[0], d[1], d[2], cs[k]}));
      ptrs.push_back(param.output[k]->template mutable_data<int8_t>(TARGET(kCUDA)));
    }
    CHECK_EQ(sum, C);
    PL_CALL(pl_split_i8_nhwc(param.x->template data<int8_t>(), C, static_cast<int>(cs.size()), ptrs.data(), cs.data(),
                             cs.data(), d[0] * d[1] * d[2], StreamOf(ctx_.get())));
  }
};

class ShuffleChannelComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::ShuffleChannelParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    const auto& d = param.X->dims();
    CHECK_EQ(d.size(), 4u);
    const int C = static_cast<int>(d[3]);
    CHECK(param.group > 0 && C % param.group == 0) << "shuffle_channel: group " << param.group << " vs C " << C;
    param.Out->Resize(d);
    PL_CALL(pl_shuffle_channel_i8_nhwc(param.X->template data<int8_t>(), C,
                                       param.Out->template mutable_data<int8_t>(TARGET(kCUDA)), C, C, param.group,
                                       d[0] * d[1] * d[2], StreamOf(ctx_.get())));
  }
};

// relu / relu6 on int8 with the input's scale kept (max(x, 0) commutes with a symmetric quantiser;
// relu6 clips at round(6 / scale)): lite/kernels/cuda/relu_compute.cu:23-33 on the integers
class ReluComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::ActivationParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    int hi = 127;
    if (param.active_type == lite_api::ActivationType::kRelu6) {
      const float q = param.Relu_clipped_coef / param.input_scale;
      hi = q >= 127.f ? 127 : static_cast<int>(q + 0.5f);
    } else {
      CHECK(param.active_type == lite_api::ActivationType::kRelu) << "int8 activation: relu / relu6 only";
    }
    param.Out->Resize(param.X->dims());
    PL_CALL(pl_relu_i8(param.X->template data<int8_t>(), param.Out->template mutable_data<int8_t>(TARGET(kCUDA)),
                       param.X->numel(), hi, StreamOf(ctx_.get())));
  }
};

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle

namespace plk = paddle::lite::kernels::cuda;
using NCHW2NHWCInt8 = plk::LayoutCompute<int8_t, true>;
using NHWC2NCHWInt8 = plk::LayoutCompute<int8_t, false>;
using NCHW2NHWCFp32 = plk::LayoutCompute<float, true>;
using NHWC2NCHWFp32 = plk::LayoutCompute<float, false>;
using PoolI8I8 = plk::PoolComputeInt8<PRECISION(kInt8)>;
using PoolI8F32 = plk::PoolComputeInt8<PRECISION(kFloat)>;
using AddI8I8 = plk::ElementwiseAddInt8<PRECISION(kInt8), paddle::lite::operators::ElementwiseParam>;
using AddI8F32 = plk::ElementwiseAddInt8<PRECISION(kFloat), paddle::lite::operators::ElementwiseParam>;
using AddActI8I8 = plk::ElementwiseAddInt8<PRECISION(kInt8), paddle::lite::operators::FusionElementwiseActivationParam>;
using AddActI8F32 = plk::ElementwiseAddInt8<PRECISION(kFloat), paddle::lite::operators::FusionElementwiseActivationParam>;

#define REG_CALIB(op__)                                                                             \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNCHW, plk::CalibComputeFp32ToInt8, fp32_to_int8)         \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kAny))}) \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kAny))})  \
      .Finalize();                                                                                  \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNCHW, plk::CalibComputeInt8ToFp32, int8_to_fp32)         \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kAny))}) \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kAny))}) \
      .Finalize();
REG_CALIB(calib)
REG_CALIB(calib_once)

#define REG_LAYOUT(op__)                                                                                       \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNCHW, NCHW2NHWCInt8, int8_nchw2nhwc)                                \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW))})          \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})           \
      .Finalize();                                                                                             \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNCHW, NHWC2NCHWInt8, int8_nhwc2nchw)                                \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})          \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW))})           \
      .Finalize();                                                                                             \
  REGISTER_LITE_KERNEL(op__, kCUDA, kFloat, kNCHW, NCHW2NHWCFp32, nchw2nhwc)                                    \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW))})         \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})          \
      .Finalize();                                                                                             \
  REGISTER_LITE_KERNEL(op__, kCUDA, kFloat, kNCHW, NHWC2NCHWFp32, nhwc2nchw)                                    \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})         \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW))})          \
      .Finalize();
REG_LAYOUT(layout)
REG_LAYOUT(layout_once)

#define REG_IOCOPY(op__)                                                                                   \
  REGISTER_LITE_KERNEL(op__, kCUDA, kAny, kAny, plk::IoCopyHostToCuda, host_to_device)                      \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny))})        \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kAny), DATALAYOUT(kAny))})         \
      .Finalize();                                                                                         \
  REGISTER_LITE_KERNEL(op__, kCUDA, kAny, kAny, plk::IoCopyCudaToHost, device_to_host)                      \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kAny), DATALAYOUT(kAny))})        \
      .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny))})         \
      .Finalize();
REG_IOCOPY(io_copy)
REG_IOCOPY(io_copy_once)

REGISTER_LITE_KERNEL(pool2d, kCUDA, kInt8, kNHWC, PoolI8I8, int8_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .Finalize();
REGISTER_LITE_KERNEL(pool2d, kCUDA, kInt8, kNHWC, PoolI8F32, fp32_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})
    .Finalize();
REGISTER_LITE_KERNEL(elementwise_add, kCUDA, kInt8, kNHWC, AddI8I8, int8_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .Finalize();
REGISTER_LITE_KERNEL(elementwise_add, kCUDA, kInt8, kNHWC, AddI8F32, fp32_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})
    .Finalize();
REGISTER_LITE_KERNEL(fusion_elementwise_add_activation, kCUDA, kInt8, kNHWC, AddActI8I8, int8_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .Finalize();
REGISTER_LITE_KERNEL(fusion_elementwise_add_activation, kCUDA, kInt8, kNHWC, AddActI8F32, fp32_out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})
    .Finalize();
#define PL_I8_NHWC LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))
REGISTER_LITE_KERNEL(concat, kCUDA, kInt8, kNHWC, plk::ConcatComputeInt8, int8_nhwc)
    .BindInput("X", {PL_I8_NHWC})
    .BindOutput("Out", {PL_I8_NHWC})
    .Finalize();
REGISTER_LITE_KERNEL(split, kCUDA, kInt8, kNHWC, plk::SplitComputeInt8, int8_nhwc)
    .BindInput("X", {PL_I8_NHWC})
    .BindOutput("Out", {PL_I8_NHWC})
    .Finalize();
REGISTER_LITE_KERNEL(shuffle_channel, kCUDA, kInt8, kNHWC, plk::ShuffleChannelComputeInt8, int8_nhwc)
    .BindInput("X", {PL_I8_NHWC})
    .BindOutput("Out", {PL_I8_NHWC})
    .Finalize();
REGISTER_LITE_KERNEL(relu, kCUDA, kInt8, kNHWC, plk::ReluComputeInt8, int8_nhwc)
    .BindInput("X", {PL_I8_NHWC})
    .BindOutput("Out", {PL_I8_NHWC})
    .Finalize();
REGISTER_LITE_KERNEL(relu6, kCUDA, kInt8, kNHWC, plk::ReluComputeInt8, int8_nhwc)
    .BindInput("X", {PL_I8_NHWC})
    .BindOutput("Out", {PL_I8_NHWC})
    .Finalize();
#undef PL_I8_NHWC
