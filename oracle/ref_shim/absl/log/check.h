// Minimal stand-in for absl/log/check.h so the reference's mjpc/spline/spline.cc compiles
// in place (abseil is not installed in this image). Test infrastructure only.
#ifndef REF_SHIM_ABSL_LOG_CHECK_H_
#define REF_SHIM_ABSL_LOG_CHECK_H_
#include <cstdlib>
#include <iostream>
namespace ref_shim {
struct CheckSink {
  bool ok;
  explicit CheckSink(bool ok_) : ok(ok_) {}
  ~CheckSink() { if (!ok) { std::cerr << std::endl; std::abort(); } }
  template <typename T> CheckSink& operator<<(const T& v) { if (!ok) std::cerr << v; return *this; }
};
}  // namespace ref_shim
#define CHECK(cond) ::ref_shim::CheckSink(static_cast<bool>(cond))
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))
#endif
