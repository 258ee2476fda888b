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
class LayoutCompute : public KernelLite<TARGET(kCUDA), PrecisionOf<T>::v, DATALAYOUT(kNCHW)> {
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
    CHECK(!param.adaptive) << "adaptive int8 pooling is not supported";
    CHECK(param.pooling_type == "max" || param.pooling_type == "avg") << "pooling_type " << param.pooling_type;
    const int N = (int)d[0], H = (int)d[1], W = (int)d[2], C = (int)d[3];
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

template <PrecisionType Ptype_out>
class ElementwiseAddInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)> {
 public:
  using param_t = operators::ElementwiseParam;
  void Run() override {
    auto& param = this->Param<param_t>();
    CHECK(param.X->dims() == param.Y->dims()) << "int8 elementwise_add needs equal shapes (residual add)";
    CHECK(param.act_type == "" || param.act_type == "relu");
    param.Out->Resize(param.X->dims());
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    void* y = out_int8 ? static_cast<void*>(param.Out->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : static_cast<void*>(param.Out->template mutable_data<float>(TARGET(kCUDA)));
    const float sa = out_int8 ? param.x_input_scale / param.output_scale : param.x_input_scale;
    const float sb = out_int8 ? param.y_input_scale / param.output_scale : param.y_input_scale;
    PL_CALL(pl_add_i8(param.X->template data<int8_t>(), param.Y->template data<int8_t>(), y, param.X->numel(), sa, sb,
                      param.act_type == "relu" ? PL_ACT_RELU : PL_ACT_NONE, 0.f, out_int8 ? PL_OUT_INT8 : PL_OUT_FP32,
                      PL_ROUND_ARM, StreamOf(ctx_.get())));
  }
};

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle

namespace plk = paddle::lite::kernels::cuda;
using paddle::lite::LiteType;
using NCHW2NHWCInt8 = plk::LayoutCompute<int8_t, true>;
using NHWC2NCHWInt8 = plk::LayoutCompute<int8_t, false>;
using NCHW2NHWCFp32 = plk::LayoutCompute<float, true>;
using NHWC2NCHWFp32 = plk::LayoutCompute<float, false>;
using PoolI8I8 = plk::PoolComputeInt8<PRECISION(kInt8)>;
using PoolI8F32 = plk::PoolComputeInt8<PRECISION(kFloat)>;
using AddI8I8 = plk::ElementwiseAddInt8<PRECISION(kInt8)>;
using AddI8F32 = plk::ElementwiseAddInt8<PRECISION(kFloat)>;

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
