// Minimal restatement of Paddle-Lite's CHECK / LOG(FATAL) convention
// (lite/utils/logging.h): a failed CHECK aborts the operation.  Like the
// reference's LITE_WITH_EXCEPTION build it throws, so a test harness can observe
// the failure; the message format "file:line Check failed: ..." is the same idea.
#pragma once
#include <sstream>
#include <stdexcept>
#include <string>

namespace paddle {
namespace lite {

class FatalStream {
 public:
  FatalStream(const char* file, int line, const char* what) { ss_ << file << ":" << line << " " << what << " "; }
  [[noreturn]] ~FatalStream() noexcept(false) { throw std::runtime_error(ss_.str()); }
  template <typename T>
  FatalStream& operator<<(const T& v) {
    ss_ << v;
    return *this;
  }

 private:
  std::ostringstream ss_;
};

struct Voidify {
  void operator&(const FatalStream&) {}
};

}  // namespace lite
}  // namespace paddle

#define LITE_FATAL(what) ::paddle::lite::FatalStream(__FILE__, __LINE__, what)
#define CHECK(cond) \
  if (cond) {       \
  } else            \
    LITE_FATAL("Check failed: " #cond)
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define LOG_FATAL LITE_FATAL("Fatal:")
// cuda_utils.h: CUDA_CALL -> LOG(FATAL) on error
#define CUDA_CALL(expr)                                                                  \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) LITE_FATAL("CUDA:") << cudaGetErrorString(e__) << " in " #expr; \
  } while (0)

