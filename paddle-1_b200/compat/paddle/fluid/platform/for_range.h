// compat: platform/for_range.h — ForRange<CPUDeviceContext> = a plain loop (reference
// platform/for_range.h:29-41).  The CUDA specialisation is not needed by the new kernels.
#pragma once
#include <cstddef>
#include "paddle/fluid/platform/device_context.h"

namespace paddle {
namespace platform {

template <typename DeviceContext>
struct ForRange {
  ForRange(const DeviceContext& dev_ctx, size_t limit);
  template <typename Function> void operator()(Function func) const;
};

template <>
struct ForRange<CPUDeviceContext> {
  ForRange(const CPUDeviceContext&, size_t limit) : limit_(limit) {}
  template <typename Function>
  void operator()(Function func) const {
    for (size_t i = 0; i < limit_; ++i) func(i);
  }
  size_t limit_;
};

}  // namespace platform
}  // namespace paddle
