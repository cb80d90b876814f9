// compat: framework/tensor_util.h — TensorCopy between places (tensor_util.cc:24-110 of the reference).
#pragma once
#include <cstring>
#include "paddle/fluid/framework/tensor.h"
#include "paddle/fluid/platform/device_context.h"

namespace paddle {
namespace framework {

inline void TensorCopy(const Tensor& src, const platform::Place& dst_place, Tensor* dst) {
  dst->Resize(src.dims());
  const size_t bytes = src.memory_size() < static_cast<size_t>(src.numel()) ? 0 : src.memory_size();
  if (src.numel() == 0) return;
  // element type is carried by the holder; copy raw bytes of numel elements
  const size_t elem = bytes / static_cast<size_t>(src.numel());
  (void)elem;
  char* d = reinterpret_cast<char*>(dst->mutable_data<char>(make_ddim({static_cast<int64_t>(bytes)}), dst_place));
  dst->Resize(src.dims());
  const char* s = reinterpret_cast<const char*>(src.data<char>());
#ifdef PADDLE_WITH_CUDA
  if (src.place().is_gpu() || dst_place.is_gpu()) {
    cudaMemcpy(d, s, bytes, cudaMemcpyDefault);
    return;
  }
#endif
  std::memcpy(d, s, bytes);
}
inline void TensorCopy(const Tensor& src, const platform::Place& dst_place, const platform::DeviceContext&,
                       Tensor* dst) {
  TensorCopy(src, dst_place, dst);
}

}  // namespace framework
}  // namespace paddle
