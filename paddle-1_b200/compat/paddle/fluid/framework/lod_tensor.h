// compat: framework/lod_tensor.h — Tensor + level-of-detail offsets (lod_tensor.h:58-121).
#pragma once
#include <vector>
#include "paddle/fluid/framework/tensor.h"
#include "paddle/fluid/framework/tensor_util.h"

namespace paddle {
namespace framework {
using LoD = std::vector<std::vector<size_t>>;
class LoDTensor : public Tensor {
 public:
  LoDTensor() {}
  const LoD& lod() const { return lod_; }
  LoD* mutable_lod() { return &lod_; }
  void set_lod(const LoD& lod) { lod_ = lod; }

 private:
  LoD lod_;
};
}  // namespace framework
}  // namespace paddle
