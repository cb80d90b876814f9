// noalloc.cu — sparse SGD for callers that cannot hand us a workspace.
//
// Who this is for: the regular entry points take a caller-provided workspace because inside a
// framework the op kernel gets temporaries from the framework's own pool.  Some callers have no
// such pool and still must not allocate: a binding that only receives the op's inputs and outputs
// (a thin cgo / ctypes / JNI shim), a process whose framework has already claimed nearly all
// device memory up front (Fluid's buddy allocator does: FLAGS_fraction_of_gpu_memory_to_use), or a
// latency-critical loop that cannot afford cudaMalloc's device-wide synchronisation.  For them the
// scratch lives in a fixed-size array inside this library's device image instead.
//
// Rules that follow from that choice:
//   * the array is B200EMB_NOALLOC_SCRATCH_BYTES (8 MiB) on every GPU that loads the library,
//     for the life of the process; a request whose workspace need exceeds it is refused with
//     B200EMB_E_SCRATCH_LIMIT and nothing is launched;
//   * every call on a device uses the same array and finds the previous call's leftovers in it;
//     the sort/reduce pipeline re-initialises each counter it relies on at the start of a call, so
//     back-to-back calls on one stream are fine, but two calls that may overlap on the GPU
//     (different streams, different threads) would scribble over each other — serialise them;
//   * the first call on each device looks the array's device address up (see below).  Make that
//     first call outside any stream capture.
#include <mutex>

#include "../../include/b200emb.h"

#include <cuda_runtime.h>

namespace {

// In host code this name is only a handle for the runtime: it is NOT a device address.  Passing
// it to a kernel would fault (or, where the GPU can reach host memory, quietly use the wrong
// bytes).  The address kernels may use comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200EMB_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device (resolved on first use).
int scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) return B200EMB_E_CUDA;
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);   // with device `dev` current
    if (e != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return B200EMB_E_CUDA;
    }
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return B200EMB_OK;
}

}  // namespace

extern "C" int b200emb_sparse_sgd_noalloc_f32(float* param, int64_t height, int64_t d,
                                              const int64_t* rows, const float* grad, int64_t n,
                                              const float* lr, unsigned flags, int32_t* status,
                                              void* stream) {
  if (n < 0 || d <= 0 || height <= 0) return B200EMB_E_INVALID;
  if (n == 0) return B200EMB_OK;
  // a plan left in somebody else's workspace cannot be "ready" here
  if (flags & B200EMB_PLAN_READY) return B200EMB_E_INVALID;
  const size_t need = b200emb_workspace_bytes(n, d, sizeof(float));
  if (need == 0) return B200EMB_E_INVALID;
  if (need > static_cast<size_t>(B200EMB_NOALLOC_SCRATCH_BYTES)) return B200EMB_E_SCRATCH_LIMIT;
  // the scratch used is the CURRENT device's: the table must live there too (a caller whose current device
  // differs from the one owning its tensors would otherwise launch across devices)
  cudaPointerAttributes attr;
  int dev = -1;
  if (cudaPointerGetAttributes(&attr, param) != cudaSuccess || cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return B200EMB_E_CUDA;
  }
  if (attr.type == cudaMemoryTypeDevice && attr.device != dev) return B200EMB_E_INVALID;
  void* scratch = nullptr;
  const int rc = scratch_for_current_device(&scratch);
  if (rc != B200EMB_OK) return rc;
  return b200emb_sparse_sgd_f32(param, height, d, rows, grad, n, lr, scratch,
                                static_cast<size_t>(B200EMB_NOALLOC_SCRATCH_BYTES), flags, status,
                                stream);
}
