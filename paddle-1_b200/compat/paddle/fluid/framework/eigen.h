// compat: framework/eigen.h — EigenVector<T>::Flatten(tensor) returning a small eager array view
// (see compat/Eigen/Dense) that supports the handful of expressions the dense branches of the
// op kernels write: out.device(dev) = a + b, a * b, s * a, p - s * g.
#pragma once
#include <Eigen/Dense>
#include "paddle/fluid/framework/tensor.h"

namespace paddle {
namespace framework {

template <typename T, int MajorType = Eigen::RowMajor, typename IndexType = Eigen::DenseIndex>
struct EigenVector {
  using Type = Eigen::Map<Eigen::Array<T, 1, Eigen::Dynamic>>;
  using ConstType = Eigen::Map<const Eigen::Array<T, 1, Eigen::Dynamic>>;
  static Type Flatten(Tensor& tensor) { return Type(tensor.data<T>(), tensor.numel()); }          // NOLINT
  static ConstType Flatten(const Tensor& tensor) { return ConstType(tensor.data<T>(), tensor.numel()); }
};

}  // namespace framework
}  // namespace paddle
