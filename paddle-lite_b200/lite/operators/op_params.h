// Parameter structs the kernels read (reference: lite/operators/op_params.h).
// Field names and meanings are the reference's; only the fields on the int8
// conv / fc / mul / matmul / calib / layout / io_copy / pool / elementwise path are kept.
// MatMulParam gains the WITH_INT8_CONFIG block (the reference struct has none,
// op_params.h:1452-1458).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "lite/core/tensor.h"
#include "lite/core/types.h"

namespace paddle {
namespace lite {
namespace operators {

#define WITH_INT8_CONFIG             \
  bool enable_int8{false};           \
  float input_scale{1.0f};           \
  std::vector<float> weight_scale{}; \
  float output_scale{1.0f};          \
  int bit_length{8};

struct ParamBase {
  virtual ~ParamBase() = default;
};

struct ActivationParam : ParamBase {
  const lite::Tensor* X{};
  lite::Tensor* Out{};
  lite_api::ActivationType active_type{lite_api::ActivationType::kIndentity};
  bool has_active{false};
  float Leaky_relu_alpha{0};
  float Relu_clipped_coef{6};
  // not in the reference's ActivationParam: the int8 relu6 kernel needs the scale of X (= scale of
  // Out) to place the clip; the precision-cast pass extension fills it in
  WITH_INT8_CONFIG
};

struct ConvParam : ParamBase {
  lite::Tensor* x{};
  lite::Tensor* filter{};
  lite::Tensor* bias{nullptr};
  lite::Tensor* residualData{nullptr};
  lite::Tensor* output{};
  std::vector<int> strides{1, 1};
  std::shared_ptr<std::vector<int>> paddings;   // {top, bottom, left, right}
  int groups{1};
  std::shared_ptr<std::vector<int>> dilations;  // {dh, dw}
  bool fuse_residual_connection{false};
  float scale_in_eltwise{1.0f};  // scale of residualData when fused (int8)
  std::string data_format{"Anylayout"};
  ActivationParam activation_param;
  WITH_INT8_CONFIG
};

struct FcParam : ParamBase {
  lite::Tensor* input{nullptr};
  lite::Tensor* w{nullptr};  // [K, N]
  lite::Tensor* bias{nullptr};
  lite::Tensor* output{nullptr};
  lite::DDim in_mat_dims;
  lite::DDim w_dims;
  int in_num_col_dims{1};
  std::string activation_type{""};
  bool padding_weights{false};
  WITH_INT8_CONFIG
};

struct MulParam : ParamBase {
  const lite::Tensor* x{};
  const lite::Tensor* y{};
  lite::Tensor* output{};
  int x_num_col_dims{1};
  int y_num_col_dims{1};
  WITH_INT8_CONFIG
};

struct MatMulParam : ParamBase {
  const lite::Tensor* X{};
  const lite::Tensor* Y{};
  lite::Tensor* Out{};
  bool transpose_X{false};
  bool transpose_Y{false};
  float alpha{1.0f};
  WITH_INT8_CONFIG
};

struct CalibParam : ParamBase {
  const lite::Tensor* input{};
  lite::Tensor* output{};
  float scale;
};

struct LayoutParam : ParamBase {
  const lite::Tensor* x{};
  lite::Tensor* y{};
  int process_type{0};
};

struct IoCopyParam : ParamBase {
  const lite::Tensor* x{};
  lite::Tensor* y{};
  int process_type{0};
};

struct PoolParam : ParamBase {
  lite::Tensor* x{};
  lite::Tensor* output{};
  std::string pooling_type{""};
  std::vector<int> ksize{};
  bool global_pooling{false};
  std::vector<int> strides{1, 1};
  std::shared_ptr<std::vector<int>> paddings;
  bool exclusive{true};
  bool adaptive{false};
  bool ceil_mode{false};
  std::string data_format{"AnyLayout"};
  WITH_INT8_CONFIG
};

struct ElementwiseParam : ParamBase {
  const lite::Tensor* X{};
  const lite::Tensor* Y{};
  lite::Tensor* Out{};
  int axis{-1};
  WITH_INT8_CONFIG
  float x_input_scale{1.0};
  float y_input_scale{1.0};
};

// lite/operators/op_params.h:678-680
struct FusionElementwiseActivationParam : public ElementwiseParam {
  std::string act_type;
};

// lite/operators/op_params.h:369-389 (ConcatParam), 590-612 (SplitParam), 258-263 (ShuffleChannelParam)
struct ConcatParam : ParamBase {
  std::vector<lite::Tensor*> x{};
  lite::Tensor* output{};
  int axis{0};  // the model's (NCHW) axis: 1 = channels
};

struct SplitParam : ParamBase {
  lite::Tensor* x{};
  std::vector<lite::Tensor*> output{};
  int axis{-1};
  int num{0};
  std::vector<int> sections;
};

struct ShuffleChannelParam : ParamBase {
  const lite::Tensor* X{};
  lite::Tensor* Out{};
  int group{1};
};

}  // namespace operators
}  // namespace lite
}  // namespace paddle
