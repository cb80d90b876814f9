// gather.cu — lookup_table forward (K1): out[k,:] = table[ids[k],:], padding rows zero.
//
// Replaces LookupTable<T,128,8,8,Padding> (operators/lookup_table_op.cu:24-50), whose launch is
// pinned to 8 blocks and scalar 4-byte loads.  Here a lane group (8/16/32 lanes, chosen from D)
// owns one row, moves it with 128-bit loads/stores, keeps kRowsInFlight independent rows in
// flight per group, and the grid covers every SM at full occupancy.  Table loads allocate in
// L1/L2 (hot Zipf rows are re-read), output stores stream past L1.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace b200emb {
namespace {

#ifndef B200EMB_GATHER_ROWS
#define B200EMB_GATHER_ROWS 4
#endif
constexpr int kRowsInFlight = B200EMB_GATHER_ROWS;

template <typename T, int LPR, int NV, bool VEC>
__global__ void __launch_bounds__(kBlock, 4)
k_gather(const T* __restrict__ table, const int64_t* __restrict__ ids, T* __restrict__ out,
         int64_t n, int64_t height, int d, int64_t padding_idx, int32_t* status,
         const int64_t* __restrict__ seg_start, T* const* __restrict__ seg_dst, int segs, SignalSpec sig) {
  pdl_wait();
  pdl_launch_dependents();
  if (segs > 0) n = min(n, seg_start[segs]);   // sharded owner side: the live count exists on the device only
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int kGroups = kBlock / LPR;
  const int l = threadIdx.x % LPR;
  const int64_t group = static_cast<int64_t>(blockIdx.x) * kGroups + threadIdx.x / LPR;
  const int64_t num_groups = static_cast<int64_t>(gridDim.x) * kGroups;
  bool bad = false;
  const int64_t step = num_groups * kRowsInFlight;
  int64_t nid[kRowsInFlight];   // ids of the batch after the current one (software prefetch)
#pragma unroll
  for (int r = 0; r < kRowsInFlight; ++r) {
    const int64_t k = group + r * num_groups;
    nid[r] = k < n ? ids[k] : 0;
  }
  for (int64_t base = group; base < n; base += step) {
    int64_t id[kRowsInFlight];
    bool live[kRowsInFlight], copy[kRowsInFlight];
#pragma unroll
    for (int r = 0; r < kRowsInFlight; ++r) {
      const int64_t k = base + r * num_groups;
      live[r] = k < n;
      id[r] = nid[r];
      const bool pad = padding_idx != -1 && id[r] == padding_idx;   // lookup_table_op.h:56
      const bool ok = id[r] >= 0 && id[r] < height;                 // lookup_table_op.h:59-60
      copy[r] = live[r] && !pad && ok;
      bad |= live[r] && !pad && !ok;
    }
#pragma unroll
    for (int r = 0; r < kRowsInFlight; ++r) {
      const int64_t k = base + step + r * num_groups;
      nid[r] = k < n ? ids[k] : 0;
    }
    for (int col0 = 0; col0 < d; col0 += LPR * NV * EPV) {
      T v[kRowsInFlight][NV][EPV];
#pragma unroll
      for (int r = 0; r < kRowsInFlight; ++r)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk) {
          const int col = col0 + (l + kk * LPR) * EPV;
          if (copy[r] && col < d) {
            load_vec<T, VEC>(table + id[r] * d + col, v[r][kk]);
          } else {
#pragma unroll
            for (int e = 0; e < EPV; ++e) v[r][kk][e] = T(0);
          }
        }
#pragma unroll
      for (int r = 0; r < kRowsInFlight; ++r)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk) {
          const int col = col0 + (l + kk * LPR) * EPV;
          if (live[r] && col < d) {
            const int64_t k = base + r * num_groups;
            T* dst = out + k * d;
            if (segs > 0) {   // scatter to per-segment destinations (peer GPUs in the sharded path)
              int g = 0;
              while (g + 1 < segs && k >= seg_start[g + 1]) ++g;
              dst = seg_dst[g] + (k - seg_start[g]) * d;
            }
            store_vec_stream<T, VEC>(dst + col, v[r][kk]);
          }
        }
    }
  }
  if (bad && status) atomicOr(status, 1);
  signal_peers(sig.flag_dst, sig.payload_dst, sig.starts, sig.epoch, sig.done_ctr, sig.world);
}


// ---- wide rows: TMA bulk row copies ---------------------------------------------------------------
// For rows of 512 B and more a lane does not touch the data at all: it asks the TMA unit to copy its row
// global -> shared (cp.async.bulk + mbarrier complete_tx) and, once the warp's batch has landed, shared ->
// global (bulk store, bulk_group).  One lane = one row, up to 32 rows in flight per warp, no registers
// spent on payload.  Padding / out-of-range rows are zero-filled in shared memory by their lane.
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(kBlock)
k_gather_bulk(const T* __restrict__ table, const int64_t* __restrict__ ids, T* __restrict__ out, int64_t n,
              int64_t height, int d, int64_t padding_idx, int32_t* status, int slots) {
  constexpr int kWarps = kBlock / 32;
  extern __shared__ __align__(128) unsigned char bulk_smem[];   // [kWarps][slots][row bytes]
  __shared__ __align__(8) uint64_t bars[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t row_bytes = static_cast<uint32_t>(d) * sizeof(T);
  if (lane == 0) mbar_init(&bars[warp], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  unsigned char* my_slot = bulk_smem + (static_cast<size_t>(warp) * slots + lane) * row_bytes;
  const int64_t warp_global = static_cast<int64_t>(blockIdx.x) * kWarps + warp;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kWarps * slots;
  uint32_t parity = 0;
  bool bad = false;
  for (int64_t base = warp_global * slots; base < n; base += stride) {
    const int64_t k = base + lane;
    const bool live = lane < slots && k < n;
    const int64_t id = live ? ids[k] : 0;
    const bool pad = padding_idx != -1 && id == padding_idx;
    const bool ok = id >= 0 && id < height;
    const bool copy = live && !pad && ok;
    bad |= live && !pad && !ok;
    const uint32_t cnt = __popc(__ballot_sync(0xffffffffu, copy));
    if (lane == 0 && cnt > 0) mbar_expect_tx(&bars[warp], cnt * row_bytes);
    __syncwarp();
    if (copy) {
      bulk_g2s(my_slot, table + id * d, row_bytes, &bars[warp]);
    } else if (live) {
      for (uint32_t b = 0; b < row_bytes; b += 16) *reinterpret_cast<int4*>(my_slot + b) = make_int4(0, 0, 0, 0);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // zero fills visible to the bulk store
    if (cnt > 0) {
      mbar_wait(&bars[warp], parity);
      parity ^= 1u;
    }
    __syncwarp();
    if (live) bulk_s2g(out + k * d, my_slot, row_bytes);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // slot may be overwritten next round
    __syncwarp();
  }
  if (bad && status) atomicOr(status, 1);
}

}  // namespace

template <typename T>
cudaError_t launch_gather(const T* table, int64_t height, int64_t d, const int64_t* ids, int64_t n,
                          int64_t padding_idx, T* out, int32_t* status, cudaStream_t s,
                          const int64_t* seg_start, T* const* seg_dst, int segs, const SignalSpec& signal) {
  if (n == 0 && signal.flag_dst == nullptr) return cudaSuccess;
  const bool aligned = is_aligned16(table) && (segs > 0 || is_aligned16(out));
  // wide rows (>= 512 B): TMA bulk row copies — opt-in (B200EMB_GATHER_BULK=1).  Measured on B200 the
  // un-pipelined bulk path loses to the 128-bit LDG/STG path, which already runs near the HBM roofline for
  // wide rows (C5, 1 KB rows: 1966 us vs 1436 us; C3, 512 B rows: 84 us vs 79 us), so LDG stays the default.
  const int64_t row_bytes = d * static_cast<int64_t>(sizeof(T));
  static const bool bulk_on = std::getenv("B200EMB_GATHER_BULK") != nullptr;
  if (bulk_on && segs == 0 && aligned && row_bytes >= 512 && row_bytes % 16 == 0 && row_bytes <= 12288) {
    constexpr int kWarps = kBlock / 32;
    const int slots = static_cast<int>(std::min<int64_t>(32, (96 * 1024) / (kWarps * row_bytes)));
    const size_t smem = static_cast<size_t>(kWarps) * slots * row_bytes;
    cudaError_t e = cudaFuncSetAttribute(k_gather_bulk<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    const int64_t want = ceil_div(n, static_cast<int64_t>(kWarps) * slots);
    const int grid = static_cast<int>(std::min<int64_t>(want, 148 * 2));
    k_gather_bulk<T><<<grid, kBlock, smem, s>>>(table, ids, out, n, height, static_cast<int>(d), padding_idx,
                                                status, slots);
    return cudaGetLastError();
  }
  cudaError_t err = cudaSuccess;
  dispatch_row_shape<T>(d, aligned, [&]<int LPR, int NV, bool VEC>() {
    constexpr int kGroups = kBlock / LPR;
    // one resident wave (4 CTAs of 256 threads per SM leaves room for the concurrent sort cluster)
    static const int cps = [] {   // resident CTAs per SM the grid aims at (tuning override, read once)
      const char* ov = std::getenv("B200EMB_GATHER_CPS");
      const int c = ov ? std::atoi(ov) : 0;
      return (c >= 1 && c <= 8) ? c : 4;
    }();
    const int64_t want = ceil_div(n, static_cast<int64_t>(kGroups));
    const int grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(want, 148 * cps)));
    err = launch_row_kernel(k_gather<T, LPR, NV, VEC>, grid, kBlock, 0, s, table, ids, out, n, height,
                            static_cast<int>(d), padding_idx, status, seg_start, seg_dst, segs, signal);
  });
  return err;
}

template cudaError_t launch_gather<float>(const float*, int64_t, int64_t, const int64_t*, int64_t,
                                          int64_t, float*, int32_t*, cudaStream_t, const int64_t*,
                                          float* const*, int, const SignalSpec&);
template cudaError_t launch_gather<double>(const double*, int64_t, int64_t, const int64_t*, int64_t,
                                           int64_t, double*, int32_t*, cudaStream_t, const int64_t*,
                                           double* const*, int, const SignalSpec&);

}  // namespace b200emb
