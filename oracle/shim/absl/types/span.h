// TEST INFRASTRUCTURE ONLY (oracle build).  Minimal stand-in for absl::Span so
// that the reference's cc/kernels/range_coder.{h,cc} compile unmodified from
// /root/reference (they include "absl/types/span.h", range_coder.h:22).
#ifndef ORACLE_SHIM_ABSL_SPAN_H_
#define ORACLE_SHIM_ABSL_SPAN_H_
#include <cstddef>
#include <initializer_list>
#include <type_traits>
#include <vector>
namespace absl {
template <typename T>
class Span {
 public:
  using size_type = std::size_t;
  using value_type = typename std::remove_cv<T>::type;
  using iterator = T*;
  Span() : p_(nullptr), n_(0) {}
  Span(T* p, size_type n) : p_(p), n_(n) {}
  template <typename V, typename = typename std::enable_if<
                            std::is_const<T>::value &&
                            std::is_same<V, value_type>::value>::type>
  Span(std::initializer_list<V> l) : p_(l.begin()), n_(l.size()) {}
  template <typename A>
  Span(const std::vector<value_type, A>& v) : p_(v.data()), n_(v.size()) {}
  template <typename A, typename U = T,
            typename = typename std::enable_if<!std::is_const<U>::value>::type>
  Span(std::vector<value_type, A>& v) : p_(v.data()), n_(v.size()) {}
  T* data() const { return p_; }
  size_type size() const { return n_; }
  bool empty() const { return n_ == 0; }
  T& operator[](size_type i) const { return p_[i]; }
  T* begin() const { return p_; }
  T* end() const { return p_ + n_; }
 private:
  T* p_;
  size_type n_;
};
}  // namespace absl
#endif
