// compat: platform/device_context.h — CPUDeviceContext (with a dummy eigen_device) and
// CUDADeviceContext exposing stream() and Wait(), cf. the reference's device_context.h:42-120.
#pragma once
#include "paddle/fluid/platform/place.h"
#ifdef PADDLE_WITH_CUDA
#include <cuda_runtime.h>
#endif

namespace Eigen { struct DefaultDevice {}; struct GpuDevice {}; }

namespace paddle {
namespace platform {

class DeviceContext {
 public:
  virtual ~DeviceContext() {}
  virtual Place GetPlace() const = 0;
  virtual void Wait() const {}
};

class CPUDeviceContext : public DeviceContext {
 public:
  CPUDeviceContext() {}
  explicit CPUDeviceContext(CPUPlace) {}
  Eigen::DefaultDevice* eigen_device() const { return &dev_; }
  Place GetPlace() const override { return CPUPlace(); }

 private:
  mutable Eigen::DefaultDevice dev_;
};

#ifdef PADDLE_WITH_CUDA
class CUDADeviceContext : public DeviceContext {
 public:
  explicit CUDADeviceContext(CUDAPlace place, cudaStream_t stream = nullptr, bool own = false)
      : place_(place), stream_(stream), own_(own) {
    if (stream_ == nullptr) {
      cudaSetDevice(place.device);
      cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking);
      own_ = true;
    }
  }
  ~CUDADeviceContext() override { if (own_ && stream_) cudaStreamDestroy(stream_); }
  cudaStream_t stream() const { return stream_; }
  Place GetPlace() const override { return place_; }
  void Wait() const override { cudaStreamSynchronize(stream_); }
  Eigen::GpuDevice* eigen_device() const { return &dev_; }

 private:
  CUDAPlace place_;
  cudaStream_t stream_;
  bool own_;
  mutable Eigen::GpuDevice dev_;
};
#endif

}  // namespace platform
}  // namespace paddle
