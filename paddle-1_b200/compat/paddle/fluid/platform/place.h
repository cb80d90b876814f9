// compat: platform/place.h — CPUPlace / CUDAPlace and a two-alternative Place, with the
// is_*_place helpers and a boost::get<> lookalike (the reference's Place is a boost::variant,
// platform/place.h:25-88).
#pragma once
#include <ostream>
#include "paddle/fluid/platform/enforce.h"

namespace paddle {
namespace platform {

struct CPUPlace {
  bool operator==(const CPUPlace&) const { return true; }
  bool operator!=(const CPUPlace&) const { return false; }
};
struct CUDAPlace {
  CUDAPlace() : device(0) {}
  explicit CUDAPlace(int d) : device(d) {}
  int GetDeviceId() const { return device; }
  bool operator==(const CUDAPlace& o) const { return device == o.device; }
  bool operator!=(const CUDAPlace& o) const { return device != o.device; }
  int device;
};
struct CUDAPinnedPlace {};

class Place {
 public:
  Place() : is_gpu_(false) {}
  Place(const CPUPlace&) : is_gpu_(false) {}             // NOLINT implicit like the variant
  Place(const CUDAPlace& p) : is_gpu_(true), gpu_(p) {}  // NOLINT
  bool is_gpu() const { return is_gpu_; }
  const CUDAPlace& gpu() const { return gpu_; }
  bool operator==(const Place& o) const { return is_gpu_ == o.is_gpu_ && (!is_gpu_ || gpu_ == o.gpu_); }
  bool operator!=(const Place& o) const { return !(*this == o); }
  bool operator<(const Place& o) const {   // CPU before GPUs, GPUs by device id (the variant's ordering)
    if (is_gpu_ != o.is_gpu_) return !is_gpu_;
    return is_gpu_ && gpu_.device < o.gpu_.device;
  }

 private:
  bool is_gpu_;
  CUDAPlace gpu_;
};

inline bool is_gpu_place(const Place& p) { return p.is_gpu(); }
inline bool is_cpu_place(const Place& p) { return !p.is_gpu(); }
inline bool places_are_same_class(const Place& a, const Place& b) { return a.is_gpu() == b.is_gpu(); }
inline std::ostream& operator<<(std::ostream& os, const Place& p) {
  if (p.is_gpu()) return os << "CUDAPlace(" << p.gpu().device << ")";
  return os << "CPUPlace";
}

}  // namespace platform
}  // namespace paddle

namespace boost {
template <typename T> T get(const ::paddle::platform::Place& p);
template <> inline ::paddle::platform::CPUPlace get< ::paddle::platform::CPUPlace>(const ::paddle::platform::Place& p) {
  PADDLE_ENFORCE(!p.is_gpu(), "boost::get<CPUPlace> on a CUDAPlace");
  return ::paddle::platform::CPUPlace();
}
template <> inline ::paddle::platform::CUDAPlace get< ::paddle::platform::CUDAPlace>(const ::paddle::platform::Place& p) {
  PADDLE_ENFORCE(p.is_gpu(), "boost::get<CUDAPlace> on a CPUPlace");
  return p.gpu();
}
}  // namespace boost
