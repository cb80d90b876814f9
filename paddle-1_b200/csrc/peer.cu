// peer.cu — set-up helpers for the fused compute+exchange kernels of the sharded path: a device
// buffer other processes' GPUs can store into (CUDA IPC), and the import side.  These run once at
// set-up, never on the step path; they are the only places in the library that allocate.
#include "../../include/b200emb.h"

#include <cuda_runtime.h>
#include <cstring>

extern "C" {

int b200emb_peer_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  if (!ptr || !handle64 || bytes == 0) return B200EMB_E_INVALID;
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return B200EMB_E_CUDA; }
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) { cudaGetLastError(); cudaFree(p); return B200EMB_E_CUDA; }
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  std::memcpy(handle64, &h, 64);
  *ptr = p;
  return B200EMB_OK;
}

int b200emb_peer_open(const unsigned char* handle64, void** ptr) {
  if (!ptr || !handle64) return B200EMB_E_INVALID;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  void* p = nullptr;
  if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
    cudaGetLastError();
    return B200EMB_E_CUDA;
  }
  *ptr = p;
  return B200EMB_OK;
}

int b200emb_peer_close(void* ptr) {
  return cudaIpcCloseMemHandle(ptr) == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA;
}

int b200emb_peer_free(void* ptr) { return cudaFree(ptr) == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA; }

}  // extern "C"
