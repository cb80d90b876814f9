// gather.cu — lookup_table forward (K1): out[k,:] = table[ids[k],:], padding rows zero.
//
// Replaces LookupTable<T,128,8,8,Padding> (operators/lookup_table_op.cu:24-50), whose launch is
// pinned to 8 blocks and scalar 4-byte loads.  Here a lane group (8/16/32 lanes, chosen from D)
// owns one row, moves it with 128-bit loads/stores, keeps kRowsInFlight independent rows in
// flight per group, and the grid covers every SM at full occupancy.  Table loads allocate in
// L1/L2 (hot Zipf rows are re-read), output stores stream past L1.
#include <algorithm>

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
         const int64_t* __restrict__ seg_start, T* const* __restrict__ seg_dst, int segs) {
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
}

}  // namespace

template <typename T>
cudaError_t launch_gather(const T* table, int64_t height, int64_t d, const int64_t* ids, int64_t n,
                          int64_t padding_idx, T* out, int32_t* status, cudaStream_t s,
                          const int64_t* seg_start, T* const* seg_dst, int segs) {
  if (n == 0) return cudaSuccess;
  const bool aligned = is_aligned16(table) && (segs > 0 || is_aligned16(out));
  cudaError_t err = cudaSuccess;
  dispatch_row_shape<T>(d, aligned, [&]<int LPR, int NV, bool VEC>() {
    constexpr int kGroups = kBlock / LPR;
    // one resident wave (4 CTAs of 256 threads per SM leaves room for the concurrent sort cluster)
    const int64_t want = ceil_div(n, static_cast<int64_t>(kGroups));
    const int grid = static_cast<int>(std::min<int64_t>(want, 148 * 4));
    k_gather<T, LPR, NV, VEC><<<grid, kBlock, 0, s>>>(table, ids, out, n, height,
                                                      static_cast<int>(d), padding_idx, status,
                                                      seg_start, seg_dst, segs);
    err = cudaGetLastError();
  });
  return err;
}

template cudaError_t launch_gather<float>(const float*, int64_t, int64_t, const int64_t*, int64_t,
                                          int64_t, float*, int32_t*, cudaStream_t, const int64_t*,
                                          float* const*, int);
template cudaError_t launch_gather<double>(const double*, int64_t, int64_t, const int64_t*, int64_t,
                                           int64_t, double*, int32_t*, cudaStream_t, const int64_t*,
                                           double* const*, int);

}  // namespace b200emb
