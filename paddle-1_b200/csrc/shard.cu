// shard.cu — the pieces the row-sharded (multi-GPU) path adds around the single-GPU kernels (K6).
//
// Sharding convention = the reference's distributed lookup table: owner(id) = id % G
// (operators/split_ids_op.h:45-49), local row = id / G, rows handed back in the requester's
// original order (operators/merge_ids_op.h:63-87).  Instead of a separate bucketing pass, ids are
// sorted once by key = owner * shard_height + local_row (KeyMap, internal.h): that one plan
//   - groups the ids by destination GPU (contiguous key ranges) and de-duplicates them
//     (k_head_count / k_head_write) so each unique row crosses NVLink once per direction,
//   - is the plan of the local MergeAdd that pre-reduces the gradient before it is sent,
//   - and its (uidx, perm) pair is the inverse map that expands the returned unique rows back to
//     the batch order (k_expand_unique = merge_ids).
#include <algorithm>

#include "common.cuh"
#include "internal.h"

namespace b200emb {
namespace {

// bounds[g] = first unique-key index owned by shard g (bounds[G] = U); local[u] = key % shard_height
// elem_bounds[g] = first SORTED POSITION owned by shard g (elem_bounds[G] = number of valid ids)
__global__ void __launch_bounds__(256)
k_shard_split(const int64_t* __restrict__ ukeys, const PlanHeader* hdr, int shards,
              uint32_t shard_height, const uint32_t* __restrict__ seg_start,
              const uint32_t* __restrict__ skey, int n, uint32_t invalid_key,
              int64_t* __restrict__ bounds, int64_t* __restrict__ elem_bounds,
              int64_t* __restrict__ local_rows) {
  const int u_total = static_cast<int>(hdr->n_unique);
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if ((bounds || elem_bounds) && tid <= shards) {
    const int64_t target = static_cast<int64_t>(tid) * shard_height;
    int lo = 0, hi = u_total;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (ukeys[mid] < target) lo = mid + 1; else hi = mid;
    }
    if (tid == shards) lo = u_total;
    if (bounds) bounds[tid] = lo;
    if (elem_bounds) {
      int64_t e;
      if (lo < u_total) {
        e = seg_start[lo];
      } else {  // past the last unique row: the valid ids end where the out-of-range keys begin
        int a = 0, b = n;
        while (a < b) {
          const int mid = (a + b) >> 1;
          if (skey[mid] < invalid_key) a = mid + 1; else b = mid;
        }
        e = a;
      }
      elem_bounds[tid] = e;
    }
  }
  if (local_rows)
    for (int u = tid; u < u_total; u += gridDim.x * blockDim.x)
      local_rows[u] = ukeys[u] % shard_height;
}

// dst[g][u - bounds[g]] = src[u] for bounds[g] <= u < bounds[g+1]; dst may be peer-GPU memory
__global__ void __launch_bounds__(256)
k_segment_scatter_i64(const int64_t* __restrict__ src, const int64_t* __restrict__ bounds,
                      int64_t* const* __restrict__ dst, int segs) {
  const int64_t total = bounds[segs];
  for (int64_t u = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; u < total;
       u += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int g = 0;
    while (g + 1 < segs && u >= bounds[g + 1]) ++g;
    dst[g][u - bounds[g]] = src[u];
  }
}

// MODE 0: out[e, :] = src[perm[e], :]   1: out[perm[e], :] = src[uidx[e], :]   2: out[perm[e], :] = src[e, :]
template <typename T, int LPR, int NV, bool VEC, int MODE>
__global__ void __launch_bounds__(kBlock, 4)
k_permute_rows(const T* __restrict__ src, const uint32_t* __restrict__ perm,
               const uint32_t* __restrict__ uidx, int n, int d, T* __restrict__ out, WaitSpec wait) {
  wait_for_flags(wait.flags, wait.epoch, wait.world, wait.status);   // rows stored by other GPUs (sharded lookup)
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int kGroups = kBlock / LPR;
  constexpr int kRows = 4;
  const int l = threadIdx.x % LPR;
  const int group = blockIdx.x * kGroups + threadIdx.x / LPR;
  const int num_groups = gridDim.x * kGroups;
  for (int base = group; base < n; base += num_groups * kRows) {
    int64_t from[kRows], to[kRows];
    bool live[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int e = base + r * num_groups;
      live[r] = e < n;
      if (MODE == 1) {
        from[r] = live[r] ? uidx[e] : 0;
        to[r] = live[r] ? perm[e] : 0;
      } else if (MODE == 2) {
        from[r] = e;
        to[r] = live[r] ? perm[e] : 0;
      } else {
        from[r] = live[r] ? perm[e] : 0;
        to[r] = e;
      }
    }
    for (int col0 = 0; col0 < d; col0 += LPR * NV * EPV) {
      T v[kRows][NV][EPV];
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          if (live[r] && col < d) {
            if (from[r] != static_cast<int64_t>(kNoKey)) {
              load_vec<T, VEC>(src + from[r] * d + col, v[r][k]);
            } else {   // an out-of-range id of the batch (no unique row): zeros, as the header promises
#pragma unroll
              for (int e = 0; e < EPV; ++e) v[r][k][e] = T(0);
            }
          }
        }
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          if (live[r] && col < d) store_vec_stream<T, VEC>(out + to[r] * d + col, v[r][k]);
        }
    }
  }
}

// inv[position in the batch] = unique index of that id: the plan's (perm, uidx) pair turned around, so that
// the expansion to batch order can walk the OUTPUT sequentially (streaming stores) instead of scattering rows
__global__ void __launch_bounds__(256)
k_inverse_map(const uint32_t* __restrict__ perm, const uint32_t* __restrict__ uidx, int n,
              uint32_t* __restrict__ inv) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) inv[perm[e]] = uidx[e];
}

}  // namespace

cudaError_t launch_inverse_map(const SortedPlan& plan, const uint32_t* uidx, int64_t n, uint32_t* inv,
                               cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, 256), 148 * 8));
  k_inverse_map<<<grid, 256, 0, s>>>(plan.perm, uidx, static_cast<int>(n), inv);
  return cudaGetLastError();
}

cudaError_t launch_shard_split(const int64_t* ukeys, const SortedPlan& plan, char* ws,
                               const Workspace& L, int shards, uint32_t shard_height, int64_t n,
                               uint32_t invalid_key, int64_t* bounds, int64_t* elem_bounds,
                               int64_t* local_rows, cudaStream_t s) {
  const int grid = static_cast<int>(std::min<int64_t>(
      std::max<int64_t>(ceil_div(std::max<int64_t>(n, shards + 1), 256), 1), 148 * 4));
  k_shard_split<<<grid, 256, 0, s>>>(ukeys, reinterpret_cast<const PlanHeader*>(ws + L.off_plan), shards,
                                     shard_height, reinterpret_cast<const uint32_t*>(ws + L.off_seg_start),
                                     plan.keys, static_cast<int>(n), invalid_key, bounds, elem_bounds,
                                     local_rows);
  return cudaGetLastError();
}

cudaError_t launch_segment_scatter_i64(const int64_t* src, int64_t n_upper, const int64_t* bounds,
                                       int64_t* const* dst, int segs, cudaStream_t s) {
  const int grid = static_cast<int>(std::min<int64_t>(std::max<int64_t>(ceil_div(n_upper, 256), 1), 148 * 8));
  k_segment_scatter_i64<<<grid, 256, 0, s>>>(src, bounds, dst, segs);
  return cudaGetLastError();
}

template <typename T>
cudaError_t launch_permute_rows(const T* src, const SortedPlan& plan, const uint32_t* uidx, int64_t n,
                                int64_t d, T* out, int mode, cudaStream_t s, const WaitSpec& wait) {
  if (n == 0) return cudaSuccess;
  const bool aligned = is_aligned16(src) && is_aligned16(out);
  cudaError_t err = cudaSuccess;
  dispatch_row_shape<T>(d, aligned, [&]<int LPR, int NV, bool VEC>() {
    constexpr int kGroups = kBlock / LPR;
    const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, kGroups), 148 * 4));
    if (mode == 1)
      k_permute_rows<T, LPR, NV, VEC, 1><<<grid, kBlock, 0, s>>>(src, plan.perm, uidx, static_cast<int>(n),
                                                                 static_cast<int>(d), out, wait);
    else if (mode == 2)
      k_permute_rows<T, LPR, NV, VEC, 2><<<grid, kBlock, 0, s>>>(src, plan.perm, uidx, static_cast<int>(n),
                                                                 static_cast<int>(d), out, wait);
    else
      k_permute_rows<T, LPR, NV, VEC, 0><<<grid, kBlock, 0, s>>>(src, plan.perm, uidx, static_cast<int>(n),
                                                                 static_cast<int>(d), out, wait);
    err = cudaGetLastError();
  });
  return err;
}

template cudaError_t launch_permute_rows<float>(const float*, const SortedPlan&, const uint32_t*, int64_t,
                                                int64_t, float*, int, cudaStream_t, const WaitSpec&);
template cudaError_t launch_permute_rows<double>(const double*, const SortedPlan&, const uint32_t*, int64_t,
                                                 int64_t, double*, int, cudaStream_t, const WaitSpec&);

}  // namespace b200emb
