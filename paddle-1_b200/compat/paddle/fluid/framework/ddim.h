// compat: framework/ddim.h — dynamic dimension vector with the helpers the op kernels call
// (make_ddim, vectorize, product, flatten_to_2d; reference ddim.h:29-130).
#pragma once
#include <cstdint>
#include <initializer_list>
#include <ostream>
#include <vector>
#include "paddle/fluid/platform/enforce.h"

namespace paddle {
namespace framework {

class DDim {
 public:
  DDim() {}
  DDim(std::initializer_list<int64_t> d) : d_(d) {}  // NOLINT
  explicit DDim(const std::vector<int64_t>& d) : d_(d) {}
  int64_t operator[](int i) const { return d_.at(static_cast<size_t>(i)); }
  int64_t& operator[](int i) { return d_.at(static_cast<size_t>(i)); }
  int size() const { return static_cast<int>(d_.size()); }
  bool operator==(const DDim& o) const { return d_ == o.d_; }
  bool operator!=(const DDim& o) const { return d_ != o.d_; }
  const std::vector<int64_t>& vec() const { return d_; }

 private:
  std::vector<int64_t> d_;
};

inline DDim make_ddim(const std::vector<int64_t>& dims) { return DDim(dims); }
inline DDim make_ddim(std::initializer_list<int64_t> dims) { return DDim(dims); }
inline DDim make_ddim(const std::vector<int>& dims) {
  return DDim(std::vector<int64_t>(dims.begin(), dims.end()));
}
inline std::vector<int64_t> vectorize(const DDim& d) { return d.vec(); }
inline int64_t product(const DDim& d) {
  int64_t p = 1;
  for (int64_t v : d.vec()) p *= v;
  return p;
}
inline int arity(const DDim& d) { return d.size(); }
inline DDim slice_ddim(const DDim& d, int begin, int end) {
  return DDim(std::vector<int64_t>(d.vec().begin() + begin, d.vec().begin() + end));
}
// [d0*...*d(k-1), dk*...]
inline DDim flatten_to_2d(const DDim& src, int num_col_dims) {
  int64_t a = 1, b = 1;
  for (int i = 0; i < src.size(); ++i) (i < num_col_dims ? a : b) *= src[i];
  return DDim({a, b});
}
inline std::ostream& operator<<(std::ostream& os, const DDim& d) {
  os << "[";
  for (int i = 0; i < d.size(); ++i) os << (i ? ", " : "") << d[i];
  return os << "]";
}

}  // namespace framework
}  // namespace paddle
