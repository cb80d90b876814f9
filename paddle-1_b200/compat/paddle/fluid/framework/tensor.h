// compat: framework/tensor.h — a dense tensor over a ref-counted allocation on a Place.
// Same surface as the reference's tensor.h:33-160 for what the embedding op kernels use:
// dims/numel/Resize, data<T>(), mutable_data<T>(place) (re-uses the holder when it is large
// enough, tensor.cc:34-74), ShareDataWith, place(), type(), IsInitialized().
#pragma once
#include <cstdlib>
#include <cstring>
#include <memory>
#include <type_traits>
#include <typeindex>
#include "paddle/fluid/framework/ddim.h"
#include "paddle/fluid/memory/memcpy.h"
#include "paddle/fluid/platform/place.h"
#ifdef PADDLE_WITH_CUDA
#include <cuda_runtime.h>
#endif

namespace paddle {
namespace framework {

class Tensor {
 public:
  struct Holder {
    void* ptr = nullptr;
    size_t bytes = 0;
    platform::Place place;
    std::type_index type = std::type_index(typeid(void));
    bool owned = true;
    ~Holder() {
      if (!ptr || !owned) return;
#ifdef PADDLE_WITH_CUDA
      if (place.is_gpu()) { cudaSetDevice(place.gpu().device); cudaFree(ptr); return; }
#endif
      std::free(ptr);
    }
  };

  Tensor() {}
  const DDim& dims() const { return dims_; }
  int64_t numel() const { return product(dims_); }
  Tensor& Resize(const DDim& d) { dims_ = d; return *this; }
  bool IsInitialized() const { return holder_ != nullptr; }
  platform::Place place() const {
    PADDLE_ENFORCE_NOT_NULL(holder_.get(), "Tensor not initialized yet when Tensor::place() is called.");
    return holder_->place;
  }
  std::type_index type() const {
    PADDLE_ENFORCE_NOT_NULL(holder_.get(), "Tensor not initialized yet when Tensor::type() is called.");
    return holder_->type;
  }

  template <typename T>
  const T* data() const {
    check_memory_size<T>();
    return reinterpret_cast<const T*>(static_cast<char*>(holder_->ptr) + offset_);
  }
  template <typename T>
  T* data() {
    check_memory_size<T>();
    return reinterpret_cast<T*>(static_cast<char*>(holder_->ptr) + offset_);
  }
  template <typename T>
  T* mutable_data(platform::Place place) {
    PADDLE_ENFORCE_GE(numel(), 0, "When calling this method, the Tensor's numel must be >= 0.");
    const size_t need = static_cast<size_t>(numel()) * sizeof(T);
    if (!holder_ || !(holder_->place == place) || holder_->bytes < need + offset_) {
      auto h = std::make_shared<Holder>();
      h->place = place;
      h->bytes = need;
      const size_t alloc = need ? need : 1;
#ifdef PADDLE_WITH_CUDA
      if (place.is_gpu()) {
        cudaSetDevice(place.gpu().device);
        PADDLE_ENFORCE(cudaMalloc(&h->ptr, alloc) == cudaSuccess, "cudaMalloc of %zu bytes failed", alloc);
      } else {
        h->ptr = std::malloc(alloc);
      }
#else
      PADDLE_ENFORCE(!place.is_gpu(), "this build has no CUDA");
      h->ptr = std::malloc(alloc);
#endif
      holder_ = h;
      offset_ = 0;
    }
    holder_->type = std::type_index(typeid(T));
    return reinterpret_cast<T*>(static_cast<char*>(holder_->ptr) + offset_);
  }
  template <typename T>
  T* mutable_data(const DDim& dims, platform::Place place) {
    Resize(dims);
    return mutable_data<T>(place);
  }
  Tensor& ShareDataWith(const Tensor& src) {
    holder_ = src.holder_;
    offset_ = src.offset_;
    dims_ = src.dims_;
    return *this;
  }
  size_t memory_size() const { return holder_ ? holder_->bytes - offset_ : 0; }
  // (compat extension) view caller-owned memory without copying; used by the harnesses only
  template <typename T>
  void ShareExternalData(T* ptr, const DDim& dims, platform::Place place) {
    auto h = std::make_shared<Holder>();
    h->ptr = const_cast<typename std::remove_const<T>::type*>(ptr);
    h->bytes = static_cast<size_t>(product(dims)) * sizeof(T);
    h->place = place;
    h->type = std::type_index(typeid(typename std::remove_const<T>::type));
    h->owned = false;
    holder_ = h;
    offset_ = 0;
    dims_ = dims;
  }

 private:
  template <typename T>
  void check_memory_size() const {
    PADDLE_ENFORCE_NOT_NULL(holder_.get(), "Tensor holds no memory. Call Tensor::mutable_data first.");
    PADDLE_ENFORCE_LE(static_cast<size_t>(numel()) * sizeof(T), memory_size(),
                      "Tensor's dims_ is out of bound.");
  }
  std::shared_ptr<Holder> holder_;
  DDim dims_;
  size_t offset_ = 0;
};

}  // namespace framework
}  // namespace paddle
