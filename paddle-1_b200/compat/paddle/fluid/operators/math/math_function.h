// compat: operators/math/math_function.h — only SetConstant (math_function_impl.h:25-30 of the
// reference: fill a tensor with one value); the BLAS half of that header is not on this path.
#pragma once
#include "paddle/fluid/framework/eigen.h"
#include "paddle/fluid/framework/tensor.h"
#include "paddle/fluid/platform/device_context.h"

namespace paddle {
namespace operators {
namespace math {

template <typename DeviceContext, typename T>
struct SetConstant {
  void operator()(const DeviceContext& context, framework::Tensor* tensor, T num);
};

template <typename T>
struct SetConstant<platform::CPUDeviceContext, T> {
  void operator()(const platform::CPUDeviceContext&, framework::Tensor* tensor, T num) {
    T* p = tensor->data<T>();
    const int64_t n = tensor->numel();
    for (int64_t i = 0; i < n; ++i) p[i] = num;
  }
};

}  // namespace math
}  // namespace operators
}  // namespace paddle
