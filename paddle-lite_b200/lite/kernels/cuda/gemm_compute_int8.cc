// fc / mul / matmul kCUDA kInt8 kernels on the tcgen05 GEMM (pl_gemm_i8).
// Shape contracts: lite/kernels/cuda/fc_compute.cu:199-353 (W is [K,N], out [M,N],
// M = prod(x.dims[:in_num_col_dims])), lite/kernels/cuda/mul_compute.h (x/y flattened by
// *_num_col_dims), lite/kernels/cuda/matmul_compute.cc:46-131 (transpose_X/Y, alpha,
// leading batch dims).  int8 semantics follow lite/kernels/arm/fc_compute.cc:83-344:
// per-output-channel (or single) weight scale, folded bias, optional relu.  mul /
// matmul int8 have no kernel on any reference target ("parity unpinned"):
// defined by analogy with fc (no bias, no activation, alpha folded into the scale).
#include "lite/core/op_registry.h"
#include "lite/kernels/cuda/int8_common.h"

namespace paddle {
namespace lite {
namespace kernels {
namespace cuda {

// B given as [K, N] row-major int8 on device -> private [N, ldk] K-major copy
static void PackKN(const int8_t* w_kn, int64_t K, int64_t N, Tensor* out, int64_t* ldk, cudaStream_t s) {
  *ldk = RoundUp(K, 16);
  out->Resize(DDim({N, *ldk}));
  int8_t* dst = out->mutable_data<int8_t>(TARGET(kCUDA));
  CUDA_CALL(cudaMemsetAsync(dst, 0, static_cast<size_t>(N * *ldk), s));
  PL_CALL(pl_transpose_i8(w_kn, static_cast<int>(N), dst, static_cast<int>(*ldk), static_cast<int>(K),
                          static_cast<int>(N), reinterpret_cast<pl_stream_t>(s)));
}

// A [M, K] dense -> pitch ldk when K % 16 != 0
static const int8_t* PadA(const int8_t* a, int64_t M, int64_t K, int64_t ldk, Tensor* scratch, cudaStream_t s) {
  if (ldk == K) return a;
  scratch->Resize(DDim({M, ldk}));
  int8_t* p = scratch->mutable_data<int8_t>(TARGET(kCUDA));
  CUDA_CALL(cudaMemcpy2DAsync(p, static_cast<size_t>(ldk), a, static_cast<size_t>(K), static_cast<size_t>(K),
                              static_cast<size_t>(M), cudaMemcpyDeviceToDevice, s));
  return p;
}

template <PrecisionType Ptype_out>
class FcComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::FcParam;
  void PrepareForRun() override {
    auto& param = this->Param<param_t>();
    auto& ctx = this->ctx_->template As<CUDAContext>();
    CHECK(param.input && param.w && param.output);
    CHECK_EQ(param.w->dims().size(), 2u);
    K_ = param.w->dims()[0];
    N_ = param.w->dims()[1];
    CHECK(param.activation_type == "" || param.activation_type == "relu")
        << "fc supports only relu fusion, got " << param.activation_type;
    epi_.Fold(param.weight_scale, static_cast<int>(N_), param.input_scale, param.output_scale,
              Ptype_out == PRECISION(kInt8), param.bias, param.activation_type == "relu" ? PL_ACT_RELU : PL_ACT_NONE, 0.f);
    PackKN(param.w->template data<int8_t>(), K_, N_, &wt_, &ldk_, ctx.exec_stream());
  }
  void Run() override {
    auto& param = this->Param<param_t>();
    auto& ctx = this->ctx_->template As<CUDAContext>();
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    const auto& xd = param.input->dims();
    const int64_t M = xd.count(0, param.in_num_col_dims);
    CHECK_EQ(xd.count(param.in_num_col_dims, static_cast<int>(xd.size())), K_) << "fc input width != W rows";
    std::vector<int64_t> od(xd.Vectorize().begin(), xd.Vectorize().begin() + param.in_num_col_dims);
    od.push_back(N_);
    param.output->Resize(od);
    void* y = out_int8 ? static_cast<void*>(param.output->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : static_cast<void*>(param.output->template mutable_data<float>(TARGET(kCUDA)));
    const int8_t* a = PadA(param.input->template data<int8_t>(), M, K_, ldk_, &a_pad_, ctx.exec_stream());
    pl_epilogue e = epi_.Make(out_int8);
    PL_CALL(pl_gemm_i8(a, static_cast<int>(ldk_), wt_.data<int8_t>(), static_cast<int>(ldk_), y, static_cast<int>(N_),
                       static_cast<int>(M), static_cast<int>(N_), static_cast<int>(K_), &e,
                       reinterpret_cast<pl_stream_t>(ctx.exec_stream())));
  }

 private:
  FoldedEpilogue epi_;
  Tensor wt_, a_pad_;
  int64_t K_{0}, N_{0}, ldk_{0};
};

template <PrecisionType Ptype_out>
class MulComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::MulParam;
  void PrepareForRun() override {
    auto& param = this->Param<param_t>();
    auto& ctx = this->ctx_->template As<CUDAContext>();
    const auto& yd = param.y->dims();
    K_ = yd.count(0, param.y_num_col_dims);
    N_ = yd.count(param.y_num_col_dims, static_cast<int>(yd.size()));
    epi_.Fold(param.weight_scale, static_cast<int>(N_), param.input_scale, param.output_scale,
              Ptype_out == PRECISION(kInt8), nullptr, PL_ACT_NONE, 0.f);
    PackKN(param.y->template data<int8_t>(), K_, N_, &wt_, &ldk_, ctx.exec_stream());
  }
  void Run() override {
    auto& param = this->Param<param_t>();
    auto& ctx = this->ctx_->template As<CUDAContext>();
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    const auto& xd = param.x->dims();
    const int64_t M = xd.count(0, param.x_num_col_dims);
    CHECK_EQ(xd.count(param.x_num_col_dims, static_cast<int>(xd.size())), K_);
    std::vector<int64_t> od(xd.Vectorize().begin(), xd.Vectorize().begin() + param.x_num_col_dims);
    const auto& yd = param.y->dims();
    for (size_t i = param.y_num_col_dims; i < yd.size(); ++i) od.push_back(yd[i]);
    param.output->Resize(od);
    void* y = out_int8 ? static_cast<void*>(param.output->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : static_cast<void*>(param.output->template mutable_data<float>(TARGET(kCUDA)));
    const int8_t* a = PadA(param.x->template data<int8_t>(), M, K_, ldk_, &a_pad_, ctx.exec_stream());
    pl_epilogue e = epi_.Make(out_int8);
    PL_CALL(pl_gemm_i8(a, static_cast<int>(ldk_), wt_.data<int8_t>(), static_cast<int>(ldk_), y, static_cast<int>(N_),
                       static_cast<int>(M), static_cast<int>(N_), static_cast<int>(K_), &e,
                       reinterpret_cast<pl_stream_t>(ctx.exec_stream())));
  }

 private:
  FoldedEpilogue epi_;
  Tensor wt_, a_pad_;
  int64_t K_{0}, N_{0}, ldk_{0};
};

// matmul: X [..., M, K] (or [..., K, M] with transpose_X), Y [K, N] / [N, K] (transpose_Y),
// or both with the same leading batch dims.  Both operands are int8 activations/weights;
// Y is re-packed in Run() unless it is 2-D (then once in PrepareForRun).
template <PrecisionType Ptype_out>
class MatMulComputeInt8 : public KernelLite<TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)> {
 public:
  using param_t = operators::MatMulParam;
  void PrepareForRun() override {
    auto& param = this->Param<param_t>();
    const auto& yd = param.Y->dims();
    CHECK_GE(yd.size(), 2u) << "int8 matmul needs a matrix Y";
    const size_t r = yd.size();
    K_ = param.transpose_Y ? yd[r - 1] : yd[r - 2];
    N_ = param.transpose_Y ? yd[r - 2] : yd[r - 1];
    // alpha scales the product: fold it into the per-column scale
    epi_.Fold(param.weight_scale, static_cast<int>(N_), param.input_scale, param.output_scale,
              Ptype_out == PRECISION(kInt8), nullptr, PL_ACT_NONE, 0.f, param.alpha);
  }
  void Run() override {
    auto& param = this->Param<param_t>();
    auto& ctx = this->ctx_->template As<CUDAContext>();
    auto s = ctx.exec_stream();
    const bool out_int8 = Ptype_out == PRECISION(kInt8);
    const auto& xd = param.X->dims();
    const auto& yd = param.Y->dims();
    CHECK_GE(xd.size(), 2u);
    const size_t rx = xd.size(), ry = yd.size();
    const int64_t M = param.transpose_X ? xd[rx - 1] : xd[rx - 2];
    const int64_t Kx = param.transpose_X ? xd[rx - 2] : xd[rx - 1];
    CHECK_EQ(Kx, K_) << "matmul inner dims differ";
    const int64_t bx = xd.count(0, static_cast<int>(rx) - 2), by = yd.count(0, static_cast<int>(ry) - 2);
    CHECK(by == 1 || by == bx) << "Y batch must be 1 or equal X batch";
    std::vector<int64_t> od(xd.Vectorize().begin(), xd.Vectorize().end() - 2);
    od.push_back(M);
    od.push_back(N_);
    param.Out->Resize(od);
    const size_t eb = out_int8 ? 1 : 4;
    char* y = out_int8 ? reinterpret_cast<char*>(param.Out->template mutable_data<int8_t>(TARGET(kCUDA)))
                       : reinterpret_cast<char*>(param.Out->template mutable_data<float>(TARGET(kCUDA)));
    const int64_t ldk = RoundUp(K_, 16);
    pl_epilogue e = epi_.Make(out_int8);
    auto st = reinterpret_cast<pl_stream_t>(s);
    const int8_t* X = param.X->template data<int8_t>();
    const int8_t* Y = param.Y->template data<int8_t>();
    // Y -> [by][N, ldk]
    yt_.Resize(DDim({by, N_, ldk}));
    int8_t* yt = yt_.mutable_data<int8_t>(TARGET(kCUDA));
    for (int64_t b = 0; b < by; ++b) {
      int8_t* dst = yt + b * N_ * ldk;
      const int8_t* src = Y + b * K_ * N_;
      if (param.transpose_Y) {
        CUDA_CALL(cudaMemcpy2DAsync(dst, static_cast<size_t>(ldk), src, static_cast<size_t>(K_),
                                    static_cast<size_t>(K_), static_cast<size_t>(N_), cudaMemcpyDeviceToDevice, s));
      } else {
        PL_CALL(pl_transpose_i8(src, static_cast<int>(N_), dst, static_cast<int>(ldk), static_cast<int>(K_),
                                static_cast<int>(N_), st));
      }
    }
    // X -> [bx*M, ldk] (transpose per batch when transpose_X, pad when K % 16 != 0)
    const int8_t* A = X;
    if (param.transpose_X || ldk != K_) {
      xa_.Resize(DDim({bx * M, ldk}));
      int8_t* xa = xa_.mutable_data<int8_t>(TARGET(kCUDA));
      for (int64_t b = 0; b < bx; ++b) {
        if (param.transpose_X)
          PL_CALL(pl_transpose_i8(X + b * M * K_, static_cast<int>(M), xa + b * M * ldk, static_cast<int>(ldk),
                                  static_cast<int>(K_), static_cast<int>(M), st));
        else
          CUDA_CALL(cudaMemcpy2DAsync(xa + b * M * ldk, static_cast<size_t>(ldk), X + b * M * K_,
                                      static_cast<size_t>(K_), static_cast<size_t>(K_), static_cast<size_t>(M),
                                      cudaMemcpyDeviceToDevice, s));
      }
      A = xa;
    }
    if (by == 1) {
      // leading dims of X fold into M: one GEMM
      PL_CALL(pl_gemm_i8(A, static_cast<int>(ldk), yt, static_cast<int>(ldk), y, static_cast<int>(N_),
                         static_cast<int>(bx * M), static_cast<int>(N_), static_cast<int>(K_), &e, st));
    } else {
      for (int64_t b = 0; b < bx; ++b)
        PL_CALL(pl_gemm_i8(A + b * M * ldk, static_cast<int>(ldk), yt + b * N_ * ldk, static_cast<int>(ldk),
                           y + b * M * N_ * eb, static_cast<int>(N_), static_cast<int>(M), static_cast<int>(N_),
                           static_cast<int>(K_), &e, st));
    }
  }

 private:
  FoldedEpilogue epi_;
  Tensor yt_, xa_;
  int64_t K_{0}, N_{0};
};

}  // namespace cuda
}  // namespace kernels
}  // namespace lite
}  // namespace paddle

namespace plk = paddle::lite::kernels::cuda;
using FcI8I8 = plk::FcComputeInt8<PRECISION(kInt8)>;
using FcI8F32 = plk::FcComputeInt8<PRECISION(kFloat)>;
using MulI8I8 = plk::MulComputeInt8<PRECISION(kInt8)>;
using MulI8F32 = plk::MulComputeInt8<PRECISION(kFloat)>;
using MatMulI8I8 = plk::MatMulComputeInt8<PRECISION(kInt8)>;
using MatMulI8F32 = plk::MatMulComputeInt8<PRECISION(kFloat)>;

REGISTER_LITE_KERNEL(fc, kCUDA, kInt8, kNCHW, FcI8I8, int8out)
    .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Bias", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})
    .BindInput("W", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .Finalize();
REGISTER_LITE_KERNEL(fc, kCUDA, kInt8, kNCHW, FcI8F32, fp32out)
    .BindInput("Input", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Bias", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})
    .BindInput("W", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})
    .Finalize();
REGISTER_LITE_KERNEL(mul, kCUDA, kInt8, kNCHW, MulI8I8, int8out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .Finalize();
REGISTER_LITE_KERNEL(mul, kCUDA, kInt8, kNCHW, MulI8F32, fp32out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})
    .Finalize();
REGISTER_LITE_KERNEL(matmul, kCUDA, kInt8, kNCHW, MatMulI8I8, int8out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .Finalize();
REGISTER_LITE_KERNEL(matmul, kCUDA, kInt8, kNCHW, MatMulI8F32, fp32out)
    .BindInput("X", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindInput("Y", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kInt8))})
    .BindOutput("Out", {LiteType::GetTensorTy(TARGET(kCUDA), PRECISION(kFloat))})
    .Finalize();
