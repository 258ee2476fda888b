#include "lite/kernels/cuda/conv_compute_int8.h"

#include "lite/core/op_registry.h"

namespace paddle {
namespace lite {
namespace kernels {
namespace cuda {

static inline void ConvOutDims(const operators::ConvParam& p, int64_t* N, int64_t* P, int64_t* Q, int64_t* K) {
  const auto& xd = p.x->dims();       // NHWC
  const auto& fd = p.filter->dims();  // OHWI
  const auto& pads = *p.paddings;
  const auto& dil = *p.dilations;
  *N = xd[0];
  *K = fd[0];
  *P = pl_conv_out_size(static_cast<int>(xd[1]), static_cast<int>(fd[1]), dil[0], pads[0], pads[1], p.strides[0]);
  *Q = pl_conv_out_size(static_cast<int>(xd[2]), static_cast<int>(fd[2]), dil[1], pads[2], pads[3], p.strides[1]);
}

template <PrecisionType Ptype_out>
void ConvComputeInt8<Ptype_out>::PrepareForRun() {
  auto& param = this->Param<param_t>();
  auto& ctx = this->ctx_->template As<CUDAContext>();
  auto stream = ctx.exec_stream();
  CHECK(param.x && param.filter && param.output);
  CHECK(param.paddings && param.paddings->size() == 4) << "paddings must be {top,bottom,left,right}";
  CHECK(param.dilations && param.dilations->size() == 2);
  CHECK_EQ(param.x->dims().size(), 4u);
  CHECK_EQ(param.filter->dims().size(), 4u);
  const bool out_int8 = Ptype_out == PRECISION(kInt8);
  const int64_t C = param.x->dims()[3];
  const int64_t K = param.filter->dims()[0], R = param.filter->dims()[1], S = param.filter->dims()[2];
  const int64_t Cg = param.filter->dims()[3];
  CHECK_EQ(Cg * param.groups, C) << "filter channels * groups != input channels";

  int act;
  float arg;
  DecodeActivation(param.activation_param, &act, &arg);
  epi_.Fold(param.weight_scale, static_cast<int>(K), param.input_scale, param.output_scale, out_int8, param.bias, act,
            arg);

  const auto& xd0 = param.x->dims();
  const pl_s2d_geometry g1 =
      (param.groups == 1 && C <= 4 && !param.residualData)
          ? pl_s2d_conv_geometry(1, static_cast<int>(xd0[1]), static_cast<int>(xd0[2]), static_cast<int>(C),
                                 static_cast<int>(R), static_cast<int>(S), param.strides[0], param.strides[1],
                                 (*param.paddings)[0], (*param.paddings)[1], (*param.paddings)[2], (*param.paddings)[3],
                                 (*param.dilations)[0], (*param.dilations)[1])
          : pl_s2d_geometry{};
  if (g1.supported && K <= 256) {
    // stride-2 thin stem (ResNet 7x7 / MobileNet 3x3): the input is re-read as 16-byte space-to-depth pixels and the
    // conv runs as a stride-1 conv over them on the tensor cores (pl_int8.h)
    path_ = kThinS2d;
    packed_filter_.Resize(DDim({static_cast<int64_t>(pl_s2d_conv_packed_weight_bytes(static_cast<int>(K), &g1))}));
    PL_CALL(pl_pack_s2d_conv_weight_i8(param.filter->data<int8_t>(), packed_filter_.mutable_data<int8_t>(TARGET(kCUDA)),
                                       static_cast<int>(K), static_cast<int>(R), static_cast<int>(S), static_cast<int>(C),
                                       &g1, reinterpret_cast<pl_stream_t>(stream)));
  } else if (param.groups == 1 && C <= 4 &&
      pl_thin_conv_supported(static_cast<int>(R), static_cast<int>(S), static_cast<int>(C), (*param.dilations)[0],
                             (*param.dilations)[1]) &&
      !param.residualData) {
    // thin first layer, direct: one 32-bit word per pixel, the kernel gathers the windows on chip
    path_ = kThinDirect;
    packed_filter_.Resize(DDim({static_cast<int64_t>(
        pl_thin_conv_packed_weight_bytes(static_cast<int>(K), static_cast<int>(R), static_cast<int>(S)))}));
    PL_CALL(pl_pack_thin_conv_weight_i8(param.filter->data<int8_t>(),
                                        packed_filter_.mutable_data<int8_t>(TARGET(kCUDA)), static_cast<int>(K),
                                        static_cast<int>(R), static_cast<int>(S), static_cast<int>(C),
                                        reinterpret_cast<pl_stream_t>(stream)));
  } else if (param.groups == 1 && C <= 4) {
    // thin first layer, other windows: one 32-bit word per pixel, explicit im2col + GEMM
    path_ = kThinIm2col;
    ldA_ = static_cast<int>(RoundUp(R * S * 4, 16));
    std::vector<int8_t> w(static_cast<size_t>(param.filter->numel()));
    DownloadBytes(param.filter, w.data(), w.size());
    std::vector<int8_t> bt(static_cast<size_t>(K) * ldA_, 0);
    for (int64_t k = 0; k < K; ++k)
      for (int64_t t = 0; t < R * S; ++t)
        for (int64_t c = 0; c < C; ++c) bt[k * ldA_ + t * 4 + c] = w[(k * R * S + t) * C + c];
    packed_filter_.Assign<int8_t, lite::DDim, TARGET(kCUDA)>(bt.data(), DDim(std::vector<int64_t>{K, ldA_}));
  } else if (param.groups == C && K == C && Cg == 1) {
    path_ = kDepthwise;
  } else {
    // dense, or 1 < groups < C (the C-ABI loops over the groups like lite/kernels/arm/conv_gemmlike.cc:84-264):
    // the filter is packed with its own channel count C / groups
    CHECK(param.groups == 1 || Cg % 16 == 0)
        << "grouped int8 conv needs 16-byte aligned channel slices (C / groups = " << Cg << ")";
    path_ = kDense;
    const size_t bytes = pl_conv_packed_weight_bytes(static_cast<int>(K), static_cast<int>(R), static_cast<int>(S),
                                                     static_cast<int>(Cg));
    packed_filter_.Resize(DDim({static_cast<int64_t>(bytes)}));
    PL_CALL(pl_pack_conv_weight_i8(param.filter->data<int8_t>(), packed_filter_.mutable_data<int8_t>(TARGET(kCUDA)),
                                   static_cast<int>(K), static_cast<int>(R), static_cast<int>(S), static_cast<int>(Cg),
                                   reinterpret_cast<pl_stream_t>(stream)));
  }
}

template <PrecisionType Ptype_out>
void ConvComputeInt8<Ptype_out>::Run() {
  auto& param = this->Param<param_t>();
  auto& ctx = this->ctx_->template As<CUDAContext>();
  auto stream = reinterpret_cast<pl_stream_t>(ctx.exec_stream());
  const bool out_int8 = Ptype_out == PRECISION(kInt8);
  int64_t N, P, Q, K;
  ConvOutDims(param, &N, &P, &Q, &K);
  // the op's InferShape works in NCHW order (lite/operators/conv_op.cc:83-95); the NHWC
  // kernel sets its own output dims like lite/kernels/cuda/conv_compute.cc:85-100
  param.output->Resize(DDim({N, P, Q, K}));
  void* y = out_int8 ? static_cast<void*>(param.output->template mutable_data<int8_t>(TARGET(kCUDA)))
                     : static_cast<void*>(param.output->template mutable_data<float>(TARGET(kCUDA)));
  const auto& xd = param.x->dims();
  const auto& fd = param.filter->dims();
  const int64_t C = xd[3];

  pl_conv_desc d{};
  d.N = static_cast<int>(N); d.H = static_cast<int>(xd[1]); d.W = static_cast<int>(xd[2]); d.C = static_cast<int>(C);
  d.ldx = static_cast<int>(C); d.K = static_cast<int>(K); d.ldy = static_cast<int>(K);
  d.R = static_cast<int>(fd[1]); d.S = static_cast<int>(fd[2]);
  d.stride_h = param.strides[0]; d.stride_w = param.strides[1];
  d.pad_t = (*param.paddings)[0]; d.pad_b = (*param.paddings)[1];
  d.pad_l = (*param.paddings)[2]; d.pad_r = (*param.paddings)[3];
  d.dil_h = (*param.dilations)[0]; d.dil_w = (*param.dilations)[1];
  d.groups = param.groups;

  pl_epilogue e = epi_.Make(out_int8);
  if (param.residualData && param.fuse_residual_connection) {
    CHECK(param.residualData->dims() == param.output->dims()) << "residualData must have the output's NHWC dims";
    e.residual = param.residualData->template data<int8_t>();
    e.ld_res = static_cast<int>(K);
    e.res_scale = out_int8 ? param.scale_in_eltwise / param.output_scale : param.scale_in_eltwise;
  }

  const int8_t* x = param.x->template data<int8_t>();
  if (path_ == kDepthwise) {
    PL_CALL(pl_dwconv2d_i8_nhwc(x, param.filter->template data<int8_t>(), y, &d, &e, stream));
    return;
  }
  const int64_t pixels = N * xd[1] * xd[2];
  if (path_ == kThinS2d) {
    const pl_s2d_geometry g = pl_s2d_conv_geometry(d.N, d.H, d.W, d.C, d.R, d.S, d.stride_h, d.stride_w, d.pad_t, d.pad_b,
                                                   d.pad_l, d.pad_r, d.dil_h, d.dil_w);
    CHECK(g.supported);
    s2d_in_.Resize(DDim({static_cast<int64_t>(g.bytes)}));
    int8_t* xs = s2d_in_.mutable_data<int8_t>(TARGET(kCUDA));
    if (g.bytes != s2d_bytes_) {  // (re)sized: the borders are zero from here on, only the interior is rewritten
      CUDA_CALL(cudaMemsetAsync(xs, 0, g.bytes, ctx.exec_stream()));
      s2d_bytes_ = g.bytes;
    }
    PL_CALL(pl_nhwc_i8_to_s2d_i8(x, static_cast<int>(C), xs, d.N, d.C, d.H, d.W, &g, stream));
    d.ldx = 4;
    PL_CALL(pl_s2d_conv2d_i8(xs, packed_filter_.data<int8_t>(), y, &d, &e, stream));
    return;
  }
  if (path_ == kThinIm2col || path_ == kThinDirect) {
    // widen the pixel pitch to one word
    x_padded_.Resize(DDim({pixels, 4}));
    int8_t* xp = x_padded_.mutable_data<int8_t>(TARGET(kCUDA));
    if (C == 4) {
      xp = const_cast<int8_t*>(x);
    } else {
      PL_CALL(pl_repitch_i8_nhwc(x, static_cast<int>(C), xp, 4, pixels, static_cast<int>(C), stream));
    }
    if (path_ == kThinDirect) {
      d.ldx = 4;
      PL_CALL(pl_thin_conv2d_i8_nhwc(xp, packed_filter_.data<int8_t>(), y, &d, &e, stream));
      return;
    }
    im2col_.Resize(DDim({N * P * Q, ldA_}));
    int8_t* A = im2col_.mutable_data<int8_t>(TARGET(kCUDA));
    PL_CALL(pl_im2col_i8_nhwc(xp, 4, A, ldA_, d.N, d.H, d.W, 4, d.R, d.S, d.stride_h, d.stride_w, d.pad_t, d.pad_l,
                              d.dil_h, d.dil_w, static_cast<int>(P), static_cast<int>(Q), stream));
    PL_CALL(pl_gemm_i8(A, ldA_, packed_filter_.data<int8_t>(), ldA_, y, d.ldy, static_cast<int>(N * P * Q),
                       static_cast<int>(K), ldA_, &e, stream));
    return;
  }
  if (C % 16 != 0) {
    // TMA needs a 16-byte pixel pitch: stage the activation once at pitch roundup(C, 16)
    const int64_t ld = RoundUp(C, 16);
    x_padded_.Resize(DDim({pixels, ld}));
    int8_t* xp = x_padded_.mutable_data<int8_t>(TARGET(kCUDA));
    PL_CALL(pl_repitch_i8_nhwc(x, static_cast<int>(C), xp, static_cast<int>(ld), pixels, static_cast<int>(C), stream));
    x = xp;
    d.ldx = static_cast<int>(ld);
  }
  PL_CALL(pl_conv2d_i8_nhwc(x, packed_filter_.data<int8_t>(), y, &d, &e, stream));
}

template class ConvComputeInt8<PRECISION(kInt8)>;
template class ConvComputeInt8<PRECISION(kFloat)>;

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle

using ConvInt8Int8 = paddle::lite::kernels::cuda::ConvComputeInt8<PRECISION(kInt8)>;
using ConvInt8Fp32 = paddle::lite::kernels::cuda::ConvComputeInt8<PRECISION(kFloat)>;

#define REGISTER_CONV_INT8(op__)                                                                              \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNHWC, ConvInt8Fp32, fp32_out)                                     \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})        \
      .BindInput("Bias", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})                           \
      .BindInput("Filter", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})       \
      .BindInput("ResidualData", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))}) \
      .BindOutput("Output", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNHWC))})     \
      .Finalize();                                                                                            \
  REGISTER_LITE_KERNEL(op__, kCUDA, kInt8, kNHWC, ConvInt8Int8, int8_out)                                     \
      .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})        \
      .BindInput("Bias", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})                           \
      .BindInput("Filter", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})       \
      .BindInput("ResidualData", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))}) \
      .BindOutput("Output", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC))})      \
      .Finalize();

REGISTER_CONV_INT8(conv2d)
REGISTER_CONV_INT8(depthwise_conv2d)
