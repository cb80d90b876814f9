// compat: memory/memcpy.h — memory::Copy overloads by (dst place, src place), cf. the reference's
// memory/memcpy.cc:22-74.  GPU copies are asynchronous on the given stream.
#pragma once
#include <cstring>
#include "paddle/fluid/platform/place.h"
#ifdef PADDLE_WITH_CUDA
#include <cuda_runtime.h>
#endif

namespace paddle {
namespace memory {

inline void Copy(platform::CPUPlace, void* dst, platform::CPUPlace, const void* src, size_t num) {
  std::memcpy(dst, src, num);
}
#ifdef PADDLE_WITH_CUDA
inline void Copy(platform::CUDAPlace, void* dst, platform::CUDAPlace, const void* src, size_t num, cudaStream_t s) {
  cudaMemcpyAsync(dst, src, num, cudaMemcpyDeviceToDevice, s);
}
inline void Copy(platform::CUDAPlace, void* dst, platform::CPUPlace, const void* src, size_t num, cudaStream_t s) {
  cudaMemcpyAsync(dst, src, num, cudaMemcpyHostToDevice, s);
}
inline void Copy(platform::CPUPlace, void* dst, platform::CUDAPlace, const void* src, size_t num, cudaStream_t s) {
  cudaMemcpyAsync(dst, src, num, cudaMemcpyDeviceToHost, s);
}
#endif

}  // namespace memory
}  // namespace paddle
