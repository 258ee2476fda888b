// ref_extra_prelude.h -- TEST INFRASTRUCTURE ONLY.  Minimal stand-ins for the lite types that the
// reference's *test-file* golden functions name, so that their text can be compiled unmodified:
//   pool_compute_ref   lite/kernels/cuda/pool_compute_test.cc:150-249 (takes operators::PoolParam)
//   matrix_mul, transpose  lite/tests/kernels/matmul_compute_test.cc:23-46
// (lite/tests/math/pool_compute_test.cc:121 `pooling_basic` cannot be used: this fork of the file carries
// unresolved merge-conflict markers and does not compile.)
// oracle/Makefile pipes  prelude + `awk` line ranges of those files + ref_extra_tail.cc  into g++;
// the reference text is never written into this repo.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace paddle {
namespace lite {
struct DDimLite {
  std::vector<int64_t> d;
  int64_t operator[](int i) const { return d[i]; }
};
struct Tensor {
  DDimLite dims_;
  void* ptr = nullptr;
  const DDimLite& dims() const { return dims_; }
  template <typename T>
  T* data() const { return static_cast<T*>(ptr); }
  template <typename T>
  T* mutable_data() { return static_cast<T*>(ptr); }
};
namespace operators {
// field names of lite/operators/op_params.h:539-559
struct PoolParam {
  Tensor* x{};
  Tensor* output{};
  std::string pooling_type{""};
  std::vector<int> ksize{};
  bool global_pooling{false};
  std::vector<int> strides{1, 1};
  std::shared_ptr<std::vector<int>> paddings;
  bool exclusive{true};
  bool adaptive{false};
  bool ceil_mode{false};
  bool use_quantizer{false};
  std::string data_format{"NCHW"};
};
}  // namespace operators
