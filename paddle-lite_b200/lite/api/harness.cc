// C window onto the C++ plugin mirror so that tests can drive the kernel classes the
// way the reference's own CUDA kernel tests do (lite/kernels/cuda/conv_compute_test.cc
// :232-298): build tensors, create the kernel through the registry (or the miniature
// static_kernel_pick), SetParam -> SetContext -> Launch(), sync, read back.
// Test / integration aid only: a Paddle-Lite build links the kernel classes directly.
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "lite/core/mir/static_kernel_pick.h"
#include "lite/core/mir/type_cast_plan.h"
#include "lite/core/op_registry.h"
#include "lite/kernels/cuda/int8_common.h"

using namespace paddle::lite;  // NOLINT
namespace ops = paddle::lite::operators;

static thread_local std::string g_err;

#define PLK_TRY try {
#define PLK_CATCH                  \
  }                                \
  catch (const std::exception& e) { \
    g_err = e.what();              \
    cudaGetLastError();            \
    return -1;                     \
  }                                \
  return 0;

static Tensor* T(void* h) { return static_cast<Tensor*>(h); }

static std::unique_ptr<KernelBase> CreateKernel(const std::string& op, TargetType t, PrecisionType p, DataLayoutType l,
                                                PrecisionType out_prec, const std::string& alias = "") {
  auto ks = KernelRegistry::Global().Create(op, t, p, l);
  for (auto& k : ks) {
    if (!alias.empty()) {
      if (k->alias() == alias) return std::move(k);
      continue;
    }
    const auto* decl = ParamTypeRegistry::Global().Find(k->key_with_alias(), k->place());
    if (!decl) continue;
    for (const char* name : {"Output", "Out"}) {
      auto it = decl->outputs.find(name);
      if (it != decl->outputs.end() && it->second.precision == out_prec) return std::move(k);
    }
  }
  LITE_FATAL("no kernel") << op << " out_prec=" << static_cast<int>(out_prec) << " alias=" << alias;
  return nullptr;
}

// A runtime program (lite/core/program.cc:265-315 RuntimeProgram::Run): instructions = kernel instances with their
// params bound, launched in order on ONE exec stream.  While a program is recording (plk_program_begin ..
// plk_program_end) every plk_run_* call appends its kernel to the program instead of launching it.
struct Program {
  std::vector<std::unique_ptr<KernelBase>> insts;
  cudaStream_t stream{nullptr};
};
static thread_local Program* g_recording = nullptr;

static void LaunchAndSync(std::unique_ptr<KernelBase>& kp, int launches) {
  if (g_recording) {
    std::unique_ptr<KernelContext> c(new KernelContext);
    c->As<CUDAContext>().SetExecStream(g_recording->stream);
    kp->SetContext(std::move(c));
    g_recording->insts.push_back(std::move(kp));
    return;
  }
  KernelBase* k = kp.get();
  std::unique_ptr<KernelContext> ctx(new KernelContext);
  auto& cuda = ctx->As<CUDAContext>();
  cudaStream_t s;
  TargetWrapperCuda::CreateStream(&s);
  cuda.SetExecStream(s);
  k->SetContext(std::move(ctx));
  try {
    for (int i = 0; i < launches; ++i) k->Launch();
    TargetWrapperCuda::StreamSync(s);
  } catch (...) {
    cudaStreamDestroy(s);
    throw;
  }
  TargetWrapperCuda::DestroyStream(s);
}

extern "C" {

const char* plk_last_error() { return g_err.c_str(); }

void* plk_tensor_create() { return new Tensor; }
void plk_tensor_destroy(void* t) { delete T(t); }

int plk_tensor_assign(void* t, int precision, int target, const int64_t* dims, int ndim, const void* host) {
  PLK_TRY
  DDim d(std::vector<int64_t>(dims, dims + ndim));
  const bool cuda = target == static_cast<int>(TARGET(kCUDA));
  switch (static_cast<PrecisionType>(precision)) {
    case PRECISION(kInt8):
      cuda ? T(t)->Assign<int8_t, paddle::lite::DDim, TARGET(kCUDA)>(static_cast<const int8_t*>(host), d)
           : T(t)->Assign<int8_t, paddle::lite::DDim, TARGET(kHost)>(static_cast<const int8_t*>(host), d);
      break;
    case PRECISION(kInt32):
      cuda ? T(t)->Assign<int32_t, paddle::lite::DDim, TARGET(kCUDA)>(static_cast<const int32_t*>(host), d)
           : T(t)->Assign<int32_t, paddle::lite::DDim, TARGET(kHost)>(static_cast<const int32_t*>(host), d);
      break;
    case PRECISION(kFloat):
      cuda ? T(t)->Assign<float, paddle::lite::DDim, TARGET(kCUDA)>(static_cast<const float*>(host), d)
           : T(t)->Assign<float, paddle::lite::DDim, TARGET(kHost)>(static_cast<const float*>(host), d);
      break;
    default: LITE_FATAL("precision") << precision;
  }
  PLK_CATCH
}
int plk_tensor_ndim(void* t) { return static_cast<int>(T(t)->dims().size()); }
void plk_tensor_dims(void* t, int64_t* out) {
  for (size_t i = 0; i < T(t)->dims().size(); ++i) out[i] = T(t)->dims()[i];
}
// declared precision of a tensor that a not-yet-run instruction will produce (a recording program picks the
// consumer's kernel from it, as the type passes do from the producer's declaration)
void plk_tensor_set_precision(void* t, int precision) { T(t)->set_precision(static_cast<PrecisionType>(precision)); }
int plk_tensor_precision(void* t) { return static_cast<int>(T(t)->precision()); }
int plk_tensor_target(void* t) { return static_cast<int>(T(t)->target()); }
int64_t plk_tensor_bytes(void* t) { return static_cast<int64_t>(T(t)->memory_size()); }
int plk_tensor_to_host(void* t, void* dst, int64_t cap) {
  PLK_TRY
  CHECK_LE(static_cast<int64_t>(T(t)->memory_size()), cap);
  TargetWrapperCuda::DeviceSync();
  paddle::lite::kernels::cuda::DownloadBytes(T(t), dst, T(t)->memory_size());
  PLK_CATCH
}

// tensor over caller-owned device memory (a torch allocation in the benches): dims + precision + the pointer
int plk_tensor_share_external(void* t, int precision, const int64_t* dims, int ndim, void* data, int64_t bytes) {
  PLK_TRY
  T(t)->Resize(std::vector<int64_t>(dims, dims + ndim));
  T(t)->set_precision(static_cast<PrecisionType>(precision));
  T(t)->ShareExternalMemory(data, static_cast<size_t>(bytes), TARGET(kCUDA));
  PLK_CATCH
}
void* plk_tensor_raw(void* t) { return const_cast<void*>(T(t)->raw_data()); }
// host <-> device copies of a tensor's bytes on a caller-provided stream (pinned host memory: asynchronous)
int plk_tensor_copy_from_host_async(void* t, const void* src, int64_t bytes, void* stream) {
  PLK_TRY
  CHECK_LE(static_cast<size_t>(bytes), T(t)->memory_size());
  CUDA_CALL(cudaMemcpyAsync(const_cast<void*>(T(t)->raw_data()), src, static_cast<size_t>(bytes), cudaMemcpyHostToDevice,
                            static_cast<cudaStream_t>(stream)));
  PLK_CATCH
}
int plk_tensor_copy_to_host_async(void* t, void* dst, int64_t bytes, void* stream) {
  PLK_TRY
  CHECK_LE(static_cast<size_t>(bytes), T(t)->memory_size());
  CUDA_CALL(cudaMemcpyAsync(dst, T(t)->raw_data(), static_cast<size_t>(bytes), cudaMemcpyDeviceToHost,
                            static_cast<cudaStream_t>(stream)));
  PLK_CATCH
}

// ---- runtime program: record plk_run_* calls, then Launch() them in order ----
void* plk_program_begin(void* stream) {
  auto* p = new Program;
  p->stream = static_cast<cudaStream_t>(stream);
  g_recording = p;
  return p;
}
int plk_program_end() {
  g_recording = nullptr;
  return 0;
}
// move every instruction to another exec stream (a CUDA-graph capture stream, another torch stream)
int plk_program_set_stream(void* prog, void* stream) {
  PLK_TRY
  auto* p = static_cast<Program*>(prog);
  p->stream = static_cast<cudaStream_t>(stream);
  for (auto& k : p->insts) k->mutable_context()->As<CUDAContext>().SetExecStream(p->stream);
  PLK_CATCH
}
int plk_program_size(void* prog) { return static_cast<int>(static_cast<Program*>(prog)->insts.size()); }
// Instruction::Run for every instruction (KernelBase::Launch: PrepareForRun on the first call, then Run)
int plk_program_run(void* prog, int sync) {
  PLK_TRY
  auto* p = static_cast<Program*>(prog);
  for (auto& k : p->insts) k->Launch();
  if (sync) TargetWrapperCuda::StreamSync(p->stream);
  PLK_CATCH
}
void plk_program_destroy(void* prog) { delete static_cast<Program*>(prog); }

// newline separated "op/alias/target/precision/layout"
int plk_list_kernels(char* buf, int cap) {
  std::string s;
  for (auto& k : KernelRegistry::Global().DebugKeys()) s += k + "\n";
  if (static_cast<int>(s.size()) + 1 > cap) return -static_cast<int>(s.size()) - 1;
  std::memcpy(buf, s.c_str(), s.size() + 1);
  return static_cast<int>(s.size());
}

// the static_kernel_pick rule: valid_places = {kCUDA/kInt8/kNHWC, kCUDA/kInt8/kNCHW, kCUDA/kFloat/kNCHW}
int plk_pick(const char* op, int enable_int8, int n_cons, const int* cons_int8, const float* cons_scale, char* out_key,
             int cap, float* out_scale) {
  PLK_TRY
  std::vector<mir::ConsumerInfo> cons(n_cons);
  for (int i = 0; i < n_cons; ++i) {
    cons[i].enable_int8 = cons_int8[i] != 0;
    cons[i].input_scale = cons_scale ? cons_scale[i] : 1.f;
  }
  std::vector<Place> places = {Place(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)),
                               Place(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)),
                               Place(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW))};
  float os = 0.f;
  auto k = mir::StaticKernelPick(op, places, enable_int8 != 0, cons, &os);
  std::string key = k ? k->SerializedKernelType() : std::string("");
  CHECK_LT(static_cast<int>(key.size()), cap);
  std::memcpy(out_key, key.c_str(), key.size() + 1);
  if (out_scale) *out_scale = os;
  PLK_CATCH
}

// KernelGrade of every candidate of `op` on the given valid places (triples target, precision, layout), best first:
// newline separated "score op/alias/target/precision/layout" (static_kernel_pick_pass.h:53-215, cc:66-84)
int plk_score_kernels(const char* op, int n_places, const int* places3, int enable_int8, int n_in, const char* const* in_args,
                      const int* in_prec, char* buf, int cap) {
  PLK_TRY
  std::vector<Place> places;
  for (int i = 0; i < n_places; ++i)
    places.emplace_back(static_cast<TargetType>(places3[3 * i]), static_cast<PrecisionType>(places3[3 * i + 1]),
                        static_cast<DataLayoutType>(places3[3 * i + 2]));
  std::map<std::string, PrecisionType> in_types;
  for (int i = 0; i < n_in; ++i) in_types[in_args[i]] = static_cast<PrecisionType>(in_prec[i]);
  std::string s;
  for (auto& c : mir::ScoreKernels(op, places, enable_int8 != 0, in_types))
    s += std::to_string(static_cast<long long>(c.score)) + " " + c.kernel->SerializedKernelType() + "\n";
  CHECK_LT(static_cast<int>(s.size()), cap);
  std::memcpy(buf, s.c_str(), s.size() + 1);
  PLK_CATCH
}

int plk_run_conv(const char* op, void* x, void* filter, void* bias, void* residual, void* out, const int* strides,
                 const int* pads, const int* dil, int groups, int act_type, float act_coef, float input_scale,
                 const float* wscale, int n_ws, float output_scale, float residual_scale, int out_int8, int launches) {
  PLK_TRY
  ops::ConvParam p;
  p.x = T(x);
  p.filter = T(filter);
  p.bias = bias ? T(bias) : nullptr;
  p.output = T(out);
  p.strides = {strides[0], strides[1]};
  p.paddings = std::make_shared<std::vector<int>>(pads, pads + 4);
  p.dilations = std::make_shared<std::vector<int>>(dil, dil + 2);
  p.groups = groups;
  if (residual) {
    p.residualData = T(residual);
    p.fuse_residual_connection = true;
    p.scale_in_eltwise = residual_scale;
  }
  if (act_type) {
    p.activation_param.has_active = true;
    p.activation_param.active_type = static_cast<paddle::lite_api::ActivationType>(act_type);
    p.activation_param.Relu_clipped_coef = act_coef;
    p.activation_param.Leaky_relu_alpha = act_coef;
  }
  p.enable_int8 = true;
  p.input_scale = input_scale;
  p.weight_scale.assign(wscale, wscale + n_ws);
  p.output_scale = output_scale;
  auto k = CreateKernel(op, TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, launches);
  PLK_CATCH
}

int plk_run_fc(void* x, void* w, void* bias, void* out, int in_num_col_dims, const char* act, float input_scale,
               const float* wscale, int n_ws, float output_scale, int out_int8, int launches) {
  PLK_TRY
  ops::FcParam p;
  p.input = T(x);
  p.w = T(w);
  p.bias = bias ? T(bias) : nullptr;
  p.output = T(out);
  p.in_num_col_dims = in_num_col_dims;
  p.activation_type = act ? act : "";
  p.enable_int8 = true;
  p.input_scale = input_scale;
  p.weight_scale.assign(wscale, wscale + n_ws);
  p.output_scale = output_scale;
  auto k = CreateKernel("fc", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, launches);
  PLK_CATCH
}

int plk_run_mul(void* x, void* y, void* out, int x_cols, int y_cols, float input_scale, const float* wscale, int n_ws,
                float output_scale, int out_int8) {
  PLK_TRY
  ops::MulParam p;
  p.x = T(x);
  p.y = T(y);
  p.output = T(out);
  p.x_num_col_dims = x_cols;
  p.y_num_col_dims = y_cols;
  p.enable_int8 = true;
  p.input_scale = input_scale;
  p.weight_scale.assign(wscale, wscale + n_ws);
  p.output_scale = output_scale;
  auto k = CreateKernel("mul", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_matmul(void* x, void* y, void* out, int tx, int ty, float alpha, float input_scale, const float* wscale,
                   int n_ws, float output_scale, int out_int8) {
  PLK_TRY
  ops::MatMulParam p;
  p.X = T(x);
  p.Y = T(y);
  p.Out = T(out);
  p.transpose_X = tx != 0;
  p.transpose_Y = ty != 0;
  p.alpha = alpha;
  p.enable_int8 = true;
  p.input_scale = input_scale;
  p.weight_scale.assign(wscale, wscale + n_ws);
  p.output_scale = output_scale;
  auto k = CreateKernel("matmul", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_calib(void* in, void* out, float scale, int to_int8) {
  PLK_TRY
  ops::CalibParam p;
  p.input = T(in);
  p.output = T(out);
  p.scale = scale;
  auto k = CreateKernel("calib", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW), PRECISION(kUnk),
                        to_int8 ? "fp32_to_int8" : "int8_to_fp32");
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_layout(void* in, void* out, int to_nhwc) {
  PLK_TRY
  ops::LayoutParam p;
  p.x = T(in);
  p.y = T(out);
  const bool i8 = T(in)->precision() == PRECISION(kInt8);
  auto k = CreateKernel("layout", TARGET(kCUDA), i8 ? PRECISION(kInt8) : PRECISION(kFloat), DATALAYOUT(kNCHW),
                        PRECISION(kUnk),
                        std::string(i8 ? "int8_" : "") + (to_nhwc ? "nchw2nhwc" : "nhwc2nchw"));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_io_copy(void* in, void* out, int h2d) {
  PLK_TRY
  ops::IoCopyParam p;
  p.x = T(in);
  p.y = T(out);
  auto k = CreateKernel("io_copy", TARGET(kCUDA), PRECISION(kAny), DATALAYOUT(kAny), PRECISION(kUnk),
                        h2d ? "host_to_device" : "device_to_host");
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_pool(void* x, void* out, const char* type, const int* ksize, const int* strides, const int* pads,
                 int global_pooling, int exclusive, int ceil_mode, float input_scale, float output_scale,
                 int out_int8) {
  PLK_TRY
  ops::PoolParam p;
  p.x = T(x);
  p.output = T(out);
  p.pooling_type = type;
  p.ksize = {ksize[0], ksize[1]};
  p.strides = {strides[0], strides[1]};
  p.paddings = std::make_shared<std::vector<int>>(pads, pads + 4);
  p.global_pooling = global_pooling != 0;
  p.exclusive = exclusive != 0;
  p.ceil_mode = (ceil_mode & 1) != 0;
  p.adaptive = (ceil_mode & 2) != 0;  // bit 1 of the flag word: PoolParam::adaptive (ksize = output size)
  p.enable_int8 = true;
  p.input_scale = input_scale;
  p.output_scale = output_scale;
  auto k = CreateKernel("pool2d", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_add(void* x, void* y, void* out, float x_scale, float y_scale, float out_scale, const char* act,
                int out_int8) {
  PLK_TRY
  ops::FusionElementwiseActivationParam p;
  p.X = T(x);
  p.Y = T(y);
  p.Out = T(out);
  p.enable_int8 = true;
  p.x_input_scale = x_scale;
  p.y_input_scale = y_scale;
  p.output_scale = out_scale;
  p.act_type = act ? act : "";
  auto k = CreateKernel("fusion_elementwise_add_activation", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC),
                        out_int8 ? PRECISION(kInt8) : PRECISION(kFloat));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

// kernel pick + io_copy / calib / layout insertion for a small graph (text form of
// mir::ParseMiniGraph); `extension` = transparent ops inherit int8,
// `reference_decls` adds the reference's own fp32 registrations as fall-backs
int plk_plan_types(const char* graph_text, int extension, int reference_decls, char* out, int cap) {
  PLK_TRY
  const mir::MiniGraph g = mir::ParseMiniGraph(graph_text ? graph_text : "");
  std::vector<mir::KernelDeclLite> decls = mir::RegisteredDecls();
  if (reference_decls)
    for (const auto& d : mir::ReferenceFp32Decls()) decls.push_back(d);
  const std::vector<Place> places = {Place(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC)),
                                     Place(TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNCHW)),
                                     Place(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW)),
                                     Place(TARGET(kHost), PRECISION(kFloat), DATALAYOUT(kNCHW))};
  const std::string text = mir::PlanToText(g, mir::PlanTypes(g, places, decls, extension != 0));
  if (static_cast<int>(text.size()) + 1 > cap) LITE_FATAL("plan text does not fit") << text.size();
  std::memcpy(out, text.c_str(), text.size() + 1);
  PLK_CATCH
}

int plk_run_concat(void* const* xs, int n, void* out, int axis) {
  PLK_TRY
  ops::ConcatParam p;
  for (int i = 0; i < n; ++i) p.x.push_back(T(xs[i]));
  p.output = T(out);
  p.axis = axis;
  auto k = CreateKernel("concat", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC), PRECISION(kInt8));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_split(void* x, void* const* outs, int n, int axis, int num, const int* sections, int n_sections) {
  PLK_TRY
  ops::SplitParam p;
  p.x = T(x);
  for (int i = 0; i < n; ++i) p.output.push_back(T(outs[i]));
  p.axis = axis;
  p.num = num;
  p.sections.assign(sections, sections + n_sections);
  auto k = CreateKernel("split", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC), PRECISION(kInt8));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_shuffle_channel(void* x, void* out, int group) {
  PLK_TRY
  ops::ShuffleChannelParam p;
  p.X = T(x);
  p.Out = T(out);
  p.group = group;
  auto k = CreateKernel("shuffle_channel", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC), PRECISION(kInt8));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

int plk_run_relu(void* x, void* out, int relu6, float clip, float input_scale) {
  PLK_TRY
  ops::ActivationParam p;
  p.X = T(x);
  p.Out = T(out);
  p.has_active = true;
  p.active_type = relu6 ? ActivationType::kRelu6 : ActivationType::kRelu;
  p.Relu_clipped_coef = clip;
  p.enable_int8 = true;
  p.input_scale = p.output_scale = input_scale;
  auto k = CreateKernel(relu6 ? "relu6" : "relu", TARGET(kCUDA), PRECISION(kInt8), DATALAYOUT(kNHWC), PRECISION(kInt8));
  k->SetParam(p);
  LaunchAndSync(k, 1);
  PLK_CATCH
}

}  // extern "C"
