// Minimal stand-in for absl/types/span.h (test infrastructure only).
#ifndef REF_SHIM_ABSL_TYPES_SPAN_H_
#define REF_SHIM_ABSL_TYPES_SPAN_H_
#include <cstddef>
#include <type_traits>
#include <vector>
namespace absl {
template <typename T>
class Span {
 public:
  using value_type = std::remove_cv_t<T>;
  using pointer = T*;
  using iterator = T*;
  Span() : p_(nullptr), n_(0) {}
  Span(T* p, std::size_t n) : p_(p), n_(n) {}
  template <typename U, typename = std::enable_if_t<std::is_const<T>::value && std::is_same<U, value_type>::value>>
  Span(const Span<U>& o) : p_(o.data()), n_(o.size()) {}
  Span(std::vector<value_type>& v) : p_(v.data()), n_(v.size()) {}
  template <typename U = T, typename = std::enable_if_t<std::is_const<U>::value>>
  Span(const std::vector<value_type>& v) : p_(v.data()), n_(v.size()) {}
  T* data() const { return p_; }
  std::size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
  T& operator[](std::size_t i) const { return p_[i]; }
  T& at(std::size_t i) const { return p_[i]; }
 private:
  T* p_;
  std::size_t n_;
};
template <typename T> Span<T> MakeSpan(T* p, std::size_t n) { return Span<T>(p, n); }
template <typename T> Span<T> MakeSpan(std::vector<T>& v) { return Span<T>(v.data(), v.size()); }
template <typename T> Span<const T> MakeConstSpan(const T* p, std::size_t n) { return Span<const T>(p, n); }
template <typename T> Span<const T> MakeConstSpan(const std::vector<T>& v) { return Span<const T>(v.data(), v.size()); }
}  // namespace absl
#endif
