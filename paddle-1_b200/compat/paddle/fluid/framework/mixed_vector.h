// compat: framework/mixed_vector.h — Vector<T>: a host std::vector with a lazily synchronised
// device mirror and dirty flags, the behaviour of the reference's mixed_vector.h:33-382
// (CUDAData uploads when the host side is newer, CUDAMutableData marks the device side
// authoritative, host accessors download when the device side is newer).
#pragma once
#include <initializer_list>
#include <vector>
#include "paddle/fluid/platform/place.h"
#ifdef PADDLE_WITH_CUDA
#include <cuda_runtime.h>
#endif

namespace paddle {
namespace framework {

template <typename T>
class Vector {
 public:
  using value_type = T;
  Vector() {}
  explicit Vector(size_t count, const T& value = T()) : cpu_(count, value) {}
  Vector(std::initializer_list<T> init) : cpu_(init) {}                 // NOLINT
  Vector(const std::vector<T>& dat) : cpu_(dat) {}                      // NOLINT implicit like the reference
  Vector(const Vector<T>& other) { other.SyncHost(); cpu_ = other.cpu_; }
  Vector<T>& operator=(const Vector<T>& other) {
    if (this != &other) { other.SyncHost(); cpu_ = other.cpu_; MarkHostNewer(); }
    return *this;
  }
  Vector<T>& operator=(const std::vector<T>& other) { cpu_ = other; MarkHostNewer(); return *this; }
  ~Vector() { FreeDevice(); }

  T& operator[](size_t i) { MutableHost(); return cpu_[i]; }
  const T& operator[](size_t i) const { SyncHost(); return cpu_[i]; }
  size_t size() const { return gpu_newer_ ? gpu_count_ : cpu_.size(); }
  bool empty() const { return size() == 0; }
  typename std::vector<T>::iterator begin() { MutableHost(); return cpu_.begin(); }
  typename std::vector<T>::iterator end() { MutableHost(); return cpu_.end(); }
  typename std::vector<T>::const_iterator begin() const { SyncHost(); return cpu_.begin(); }
  typename std::vector<T>::const_iterator end() const { SyncHost(); return cpu_.end(); }
  T* data() { MutableHost(); return cpu_.data(); }
  const T* data() const { SyncHost(); return cpu_.data(); }
  T& front() { MutableHost(); return cpu_.front(); }
  T& back() { MutableHost(); return cpu_.back(); }
  void push_back(T elem) { MutableHost(); cpu_.push_back(elem); }
  void reserve(size_t n) { cpu_.reserve(n); }
  void resize(size_t n) { MutableHost(); cpu_.resize(n); }
  void clear() { cpu_.clear(); MarkHostNewer(); }
  template <typename It>
  void assign(It b, It e) { cpu_.assign(b, e); MarkHostNewer(); }
  template <typename It>
  void Extend(It b, It e) { MutableHost(); cpu_.insert(cpu_.end(), b, e); }
  bool operator==(const Vector<T>& o) const { SyncHost(); o.SyncHost(); return cpu_ == o.cpu_; }
  operator std::vector<T>() const { SyncHost(); return cpu_; }          // NOLINT

#ifdef PADDLE_WITH_CUDA
  // device pointer valid for reads; uploads (and waits) when the host side is newer
  const T* CUDAData(platform::Place place) const {
    PADDLE_ENFORCE(place.is_gpu(), "CUDAData() needs a CUDAPlace");
    if (!gpu_newer_ && !gpu_valid_) Upload(place);
    return gpu_;
  }
  // device pointer for writes of the current size; the device side becomes authoritative
  T* CUDAMutableData(platform::Place place) {
    PADDLE_ENFORCE(place.is_gpu(), "CUDAMutableData() needs a CUDAPlace");
    if (!gpu_newer_ && !gpu_valid_) Upload(place);
    gpu_newer_ = true;
    gpu_count_ = cpu_.size();
    return gpu_;
  }
#endif

 private:
  void MarkHostNewer() { gpu_valid_ = false; gpu_newer_ = false; }
  void MutableHost() { SyncHost(); MarkHostNewer(); }
  void SyncHost() const {
#ifdef PADDLE_WITH_CUDA
    if (gpu_newer_) {
      cpu_.resize(gpu_count_);
      if (gpu_count_) cudaMemcpy(cpu_.data(), gpu_, gpu_count_ * sizeof(T), cudaMemcpyDeviceToHost);
      gpu_newer_ = false;
      gpu_valid_ = true;
    }
#endif
  }
  void FreeDevice() {
#ifdef PADDLE_WITH_CUDA
    if (gpu_) cudaFree(gpu_);
    gpu_ = nullptr;
    gpu_cap_ = 0;
#endif
  }
#ifdef PADDLE_WITH_CUDA
  void Upload(platform::Place place) const {
    cudaSetDevice(place.gpu().device);
    if (gpu_cap_ < cpu_.size() || gpu_ == nullptr) {
      if (gpu_) cudaFree(gpu_);
      gpu_cap_ = cpu_.size() ? cpu_.size() : 1;
      cudaMalloc(reinterpret_cast<void**>(&gpu_), gpu_cap_ * sizeof(T));
    }
    if (!cpu_.empty()) cudaMemcpy(gpu_, cpu_.data(), cpu_.size() * sizeof(T), cudaMemcpyHostToDevice);
    gpu_valid_ = true;
  }
#endif
  mutable std::vector<T> cpu_;
  mutable T* gpu_ = nullptr;
  mutable size_t gpu_cap_ = 0;
  mutable size_t gpu_count_ = 0;
  mutable bool gpu_valid_ = false;   // device mirror equals the host vector
  mutable bool gpu_newer_ = false;   // device mirror is authoritative
};

}  // namespace framework
}  // namespace paddle
