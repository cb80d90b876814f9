// peer.cu — set-up helpers for the fused compute+exchange kernels of the sharded path: a device
// buffer other processes' GPUs can store into (CUDA IPC), and the import side.  These run once at
// set-up, never on the step path; they are the only places in the library that allocate.
#include "../../include/b200emb.h"

#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>

namespace {
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// one thread per rank: announce to it, then wait for it
__global__ void k_peer_barrier(unsigned long long* const* __restrict__ flags, int world, int rank,
                               unsigned long long epoch, int32_t* status) {
  const int r = threadIdx.x;
  if (r >= world) return;
  unsigned long long* dst = flags[r] + rank;
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
  const unsigned long long* src = flags[rank] + r;
  const unsigned long long t0 = global_ns();
  unsigned long long v;
  for (;;) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
    if (v >= epoch) break;
    if (global_ns() - t0 > 20000000000ull) {   // 20 s: a rank is gone; do not hang the device
      if (status) atomicOr(status, B200EMB_STATUS_BARRIER_TIMEOUT);
      break;
    }
    __nanosleep(64);
  }
}
// all-gather of one small int64 row per rank over peer memory, with its own barrier: thread r stores this
// rank's row into slot [rank] of rank r's table, publishes `epoch` in rank r's flag array (release), then
// waits for rank r's flag in the local array (acquire) — after the kernel the local table holds every row.
__global__ void k_peer_allgather(const int64_t* __restrict__ row, int cols, int64_t* const* __restrict__ tables,
                                 unsigned long long* const* __restrict__ flags, int world, int rank,
                                 unsigned long long epoch, int32_t* status) {
  const int r = threadIdx.x;
  if (r >= world) return;
  int64_t* dst = tables[r] + static_cast<size_t>(rank) * cols;
  for (int c = 0; c < cols; ++c) dst[c] = row[c];
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flags[r] + rank), "l"(epoch) : "memory");
  const unsigned long long* src = flags[rank] + r;
  const unsigned long long t0 = global_ns();
  unsigned long long v;
  for (;;) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
    if (v >= epoch) break;
    if (global_ns() - t0 > 20000000000ull) {
      if (status) atomicOr(status, B200EMB_STATUS_BARRIER_TIMEOUT);
      break;
    }
    __nanosleep(64);
  }
}

// one thread per rank / owner; world <= 1024
__global__ void k_exchange_maps(const int64_t* __restrict__ tab, int world, int cols, int rank,
                                const int64_t* __restrict__ pi, const int64_t* __restrict__ pu,
                                const int64_t* __restrict__ pg, int64_t row_bytes, int64_t* __restrict__ out) {
  const int g = threadIdx.x;
  const int w1 = world + 1;
  if (g < world) {
    int64_t before = 0;
    for (int r = 0; r < rank; ++r) before += tab[r * cols + g + 1] - tab[r * cols + g];
    out[0 * w1 + g] = pi[g] + 8 * before;
    out[3 * w1 + g] = pg[g] + row_bytes * before;
    out[2 * w1 + g] = pu[g] + row_bytes * tab[g * cols + rank];
  }
  if (g <= world) {
    int64_t start = 0;
    for (int r = 0; r < g; ++r) start += tab[r * cols + rank + 1] - tab[r * cols + rank];
    out[1 * w1 + g] = start;
  }
}
}  // namespace

extern "C" {

int b200emb_exchange_maps(const int64_t* split_table, int world, int cols, int rank,
                          const int64_t* peer_i_ptrs, const int64_t* peer_u_ptrs,
                          const int64_t* peer_g_ptrs, int64_t row_bytes, int64_t* out, void* stream) {
  if (!split_table || !peer_i_ptrs || !peer_u_ptrs || !peer_g_ptrs || !out || world < 1 || world > 1023 ||
      cols < world + 1 || rank < 0 || rank >= world)
    return B200EMB_E_INVALID;
  k_exchange_maps<<<1, ((world + 1 + 31) / 32) * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      split_table, world, cols, rank, peer_i_ptrs, peer_u_ptrs, peer_g_ptrs, row_bytes, out);
  return cudaGetLastError() == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA;
}

int b200emb_peer_allgather_i64(const int64_t* row, int cols, void* const* table_ptrs,
                               const int64_t* local_table, void* const* flag_ptrs, int world, int rank,
                               uint64_t epoch, int64_t* host_copy, int32_t* status, void* stream) {
  if (!row || !table_ptrs || !flag_ptrs || cols < 1 || cols > 4096 || world < 1 || world > 1024 || rank < 0 ||
      rank >= world || (host_copy && !local_table))
    return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  k_peer_allgather<<<1, ((world + 31) / 32) * 32, 0, s>>>(
      row, cols, reinterpret_cast<int64_t* const*>(table_ptrs),
      reinterpret_cast<unsigned long long* const*>(flag_ptrs), world, rank, epoch, status);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && host_copy)
    e = cudaMemcpyAsync(host_copy, local_table, sizeof(int64_t) * static_cast<size_t>(world) * cols,
                        cudaMemcpyDeviceToHost, s);
  return e == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA;
}

int b200emb_peer_barrier(void* const* flag_ptrs, int world, int rank, uint64_t epoch,
                         int32_t* status, void* stream) {
  if (!flag_ptrs || world < 1 || world > 1024 || rank < 0 || rank >= world) return B200EMB_E_INVALID;
  k_peer_barrier<<<1, ((world + 31) / 32) * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<unsigned long long* const*>(flag_ptrs), world, rank, epoch, status);
  return cudaGetLastError() == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA;
}

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
