// sort.cu — deterministic sort-by-id ("plan") for SelectedRows rows, sm_100a.
//
// Replaces the host-side std::set + per-row linear scan of the reference's MergeAdd
// (operators/math/selected_rows_functor.cc:193-228, .cu:235-297) with a stable LSD radix
// sort of (id -> original position) done entirely on the device:
//   k_prepare      int64 ids -> u32 keys (+ range check) and the global digit histograms of
//                  every pass in one sweep;
//   k_radix_pass   one launch per 8-bit digit: warp-level match ranking (stable), tile
//                  aggregates published to a chained-scan state array, batched decoupled
//                  look-back, scatter;
//   k_head_scan    run-length heads of the sorted keys -> unique index of every element,
//                  first position of every unique row, ascending unique rows (= std::set
//                  order), U.
// Only ceil(log2(height+1)) key bits are sorted.  The sort moves 8 B per id per pass, which is
// noise next to the 4*D B of row data each id drags through HBM; launch count and look-back
// latency are what matter, hence the 1+passes launch structure.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace b200emb {

namespace {
constexpr uint32_t kFlagAgg = 1u << 30;
constexpr uint32_t kFlagIncl = 2u << 30;
constexpr uint32_t kValMask = (1u << 30) - 1;
constexpr int kLookBatch = 16;

// Walk back over predecessor tiles' published words (one counter per tile, `stride` words
// apart) until an inclusive prefix is met.  Loads are issued kLookBatch at a time so the walk
// costs one L2 round trip per batch rather than per tile.
__device__ __forceinline__ uint32_t look_back(const uint32_t* state, int tile, size_t stride) {
  uint32_t excl = 0;
  int t = tile - 1;
  while (t >= 0) {
    uint32_t v[kLookBatch];
#pragma unroll
    for (int i = 0; i < kLookBatch; ++i) {
      const int tt = t - i;
      v[i] = tt >= 0 ? ld_relaxed(state + static_cast<size_t>(tt) * stride) : kFlagIncl;
    }
    int adv = 0;
    bool fin = false, stall = false;
#pragma unroll
    for (int i = 0; i < kLookBatch; ++i) {
      if (!fin && !stall) {
        const uint32_t f = v[i] >> 30;
        if (f == 0) {
          stall = true;
        } else {
          excl += v[i] & kValMask;
          ++adv;
          fin = (f == 2);
        }
      }
    }
    if (fin) break;
    t -= adv;
  }
  return excl;
}

__global__ void __launch_bounds__(256)
k_prepare(const int64_t* __restrict__ rows, int n, uint32_t height, int passes,
          uint32_t* __restrict__ keys, uint32_t* __restrict__ ghist, int32_t* status) {
  __shared__ uint32_t sh[kMaxPasses * kRadix];
  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < passes * kRadix; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  bool bad = false;
  const int stride = gridDim.x * blockDim.x;
  for (int base = blockIdx.x * blockDim.x; base < n; base += stride) {
    const int i = base + tid;
    const bool valid = i < n;
    uint32_t key = kNoKey;
    if (valid) {
      const int64_t r = rows[i];
      const bool ok = r >= 0 && r < static_cast<int64_t>(height);
      key = ok ? static_cast<uint32_t>(r) : height;  // out-of-range rows sort last
      bad |= !ok;
      keys[i] = key;
    }
    for (int p = 0; p < passes; ++p) {
      const uint32_t digit = valid ? ((key >> (p * kRadixBits)) & (kRadix - 1)) : 0u;
      const uint32_t peers = match_digit<kRadixBits>(digit, valid);
      if (valid && lane == __ffs(peers) - 1) atomicAdd(&sh[p * kRadix + digit], __popc(peers));
    }
  }
  __syncthreads();
  for (int i = tid; i < passes * kRadix; i += blockDim.x)
    if (sh[i]) atomicAdd(&ghist[i], sh[i]);
  if (bad && status) atomicOr(status, 2);
}

template <bool HAS_PERM_IN>
__global__ void __launch_bounds__(kSortThreads)
k_radix_pass(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ perm_in,
             uint32_t* __restrict__ keys_out, uint32_t* __restrict__ perm_out, int n, int shift,
             const uint32_t* __restrict__ ghist, uint32_t* ticket, uint32_t* state) {
  constexpr int kWarps = kSortThreads / 32;
  __shared__ uint32_t s_cnt[kWarps][kRadix];
  __shared__ uint32_t s_base[kRadix];
  __shared__ uint32_t s_warp_sum[kWarps];
  __shared__ int s_tile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = static_cast<int>(atomicAdd(ticket, 1u));
#pragma unroll
  for (int w = 0; w < kWarps; ++w) s_cnt[w][tid] = 0;
  __syncthreads();
  const int tile = s_tile;
  const int warp_start = tile * kSortTile + warp * (32 * kSortItems);

  uint32_t key[kSortItems], pm[kSortItems], rank[kSortItems];
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int idx = warp_start + j * 32 + lane;
    const bool valid = idx < n;
    key[j] = valid ? keys_in[idx] : kNoKey;
    pm[j] = valid ? (HAS_PERM_IN ? perm_in[idx] : static_cast<uint32_t>(idx)) : 0u;
  }
  // stable rank inside the warp: items are ordered (j, lane) == ascending idx
  const uint32_t lt = lanemask_lt();
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const bool valid = (warp_start + j * 32 + lane) < n;
    const uint32_t digit = valid ? ((key[j] >> shift) & (kRadix - 1)) : 0u;
    const uint32_t peers = match_digit<kRadixBits>(digit, valid);
    const int leader = valid ? __ffs(peers) - 1 : lane;
    uint32_t old = 0;
    if (valid && lane == leader) {
      old = s_cnt[warp][digit];
      s_cnt[warp][digit] = old + __popc(peers);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[j] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  // bin `tid`: exclusive offsets of each warp inside the tile, tile total
  uint32_t total = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) {
    const uint32_t c = s_cnt[w][tid];
    s_cnt[w][tid] = total;
    total += c;
  }
  uint32_t* my_state = state + static_cast<size_t>(tile) * kRadix + tid;
  uint32_t excl = 0;
  if (tile == 0) {
    st_relaxed(my_state, kFlagIncl | total);
  } else {
    st_relaxed(my_state, kFlagAgg | total);
    excl = look_back(state + tid, tile, kRadix);
    st_relaxed(my_state, kFlagIncl | (excl + total));
  }
  // exclusive scan of the global histogram over bins
  uint32_t g = ghist[tid];
  uint32_t incl = g;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_warp_sum[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) warp_off += (w < warp) ? s_warp_sum[w] : 0u;
  s_base[tid] = warp_off + incl - g + excl;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int idx = warp_start + j * 32 + lane;
    if (idx < n) {
      const uint32_t digit = (key[j] >> shift) & (kRadix - 1);
      const uint32_t pos = s_base[digit] + s_cnt[warp][digit] + rank[j];
      keys_out[pos] = key[j];
      perm_out[pos] = pm[j];
    }
  }
}

__global__ void __launch_bounds__(256)
k_head_scan(const uint32_t* __restrict__ skey, int n, uint32_t invalid_key, uint32_t* ticket,
            uint32_t* state, uint32_t* __restrict__ uidx, uint32_t* __restrict__ seg_start,
            int64_t* __restrict__ out_rows, PlanHeader* hdr, int64_t* n_unique_out,
            int num_tiles) {
  constexpr int kItems = kScanTile / 256;
  __shared__ uint32_t s_warp_sum[8];
  __shared__ uint32_t s_tile_excl;
  __shared__ int s_tile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = static_cast<int>(atomicAdd(ticket, 1u));
  __syncthreads();
  const int tile = s_tile;
  const int idx0 = tile * kScanTile + tid * kItems;
  uint32_t k[kItems];
  uint32_t prev = (idx0 > 0 && idx0 - 1 < n) ? skey[idx0 - 1] : kNoKey;
  uint32_t cnt = 0;
  uint32_t headmask = 0;
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const int idx = idx0 + j;
    k[j] = idx < n ? skey[idx] : kNoKey;
    const bool head = idx < n && k[j] != prev && k[j] != invalid_key;
    headmask |= head ? (1u << j) : 0u;
    cnt += head;
    prev = k[j];
  }
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_warp_sum[warp] = incl;
  __syncthreads();
  uint32_t warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    warp_off += (w < warp) ? s_warp_sum[w] : 0u;
    block_total += s_warp_sum[w];
  }
  if (tid == 0) {
    uint32_t excl = 0;
    if (tile == 0) {
      st_relaxed(state, kFlagIncl | block_total);
    } else {
      st_relaxed(state + tile, kFlagAgg | block_total);
      excl = look_back(state, tile, 1);
      st_relaxed(state + tile, kFlagIncl | (excl + block_total));
    }
    s_tile_excl = excl;
    if (tile == num_tiles - 1) {
      hdr->n_unique = excl + block_total;
      if (n_unique_out) *n_unique_out = static_cast<int64_t>(excl + block_total);
    }
  }
  __syncthreads();
  uint32_t run = s_tile_excl + warp_off + incl - cnt;  // heads before this thread's items
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const int idx = idx0 + j;
    if (idx < n) {
      const bool head = (headmask >> j) & 1u;
      run += head;
      const uint32_t u = run - 1u;  // (wraps for leading invalid-only data; never used)
      uidx[idx] = u;
      if (head) {
        seg_start[u] = static_cast<uint32_t>(idx);
        if (out_rows) out_rows[u] = static_cast<int64_t>(k[j]);
      }
    }
  }
}

int key_bits(int64_t height) {  // keys span [0, height] (height = the invalid sentinel)
  int b = 1;
  while ((1ull << b) <= static_cast<unsigned long long>(height)) ++b;
  return b;
}
int sort_passes(int64_t height) { return static_cast<int>(ceil_div(key_bits(height), kRadixBits)); }

}  // namespace

Workspace make_workspace(int64_t n, int64_t d, int64_t elem_size) {
  Workspace L{};
  if (n < 1) n = 1;
  L.sort_tiles = ceil_div(n, kSortTile);
  L.scan_tiles = ceil_div(n, kScanTile);
  L.max_ctas = ceil_div(n, static_cast<int64_t>(kMinGroupsPerCta) * kMinChunk) + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  L.off_hist = take(sizeof(uint32_t) * kMaxPasses * kRadix);
  L.off_tickets = take(sizeof(uint32_t) * 8);
  L.off_lookback = take(sizeof(uint32_t) * kMaxPasses * L.sort_tiles * kRadix);
  L.off_scan_state = take(sizeof(uint32_t) * L.scan_tiles);
  L.zero_bytes = off;
  L.off_plan = take(sizeof(PlanHeader));
  for (int b = 0; b < 2; ++b) L.off_keys[b] = take(sizeof(uint32_t) * n);
  for (int b = 0; b < 2; ++b) L.off_perm[b] = take(sizeof(uint32_t) * n);
  L.off_uidx = take(sizeof(uint32_t) * n);
  L.off_seg_start = take(sizeof(uint32_t) * (n + 1));
  L.off_cta_flags = take(sizeof(uint32_t) * L.max_ctas);
  L.off_cta_tail_key = take(sizeof(uint32_t) * L.max_ctas);
  L.off_cta_tail_eref = take(sizeof(uint32_t) * L.max_ctas);
  L.off_cta_head = take(static_cast<size_t>(elem_size) * L.max_ctas * d);
  L.off_cta_tail = take(static_cast<size_t>(elem_size) * L.max_ctas * d);
  L.total = off;
  return L;
}

SortedPlan plan_view(int64_t height, char* ws, const Workspace& L) {
  SortedPlan p;
  p.passes = sort_passes(height);
  p.keys = reinterpret_cast<const uint32_t*>(ws + L.off_keys[p.passes & 1]);
  p.perm = reinterpret_cast<const uint32_t*>(ws + L.off_perm[p.passes & 1]);
  return p;
}

cudaError_t launch_plan(const int64_t* rows, int64_t n, int64_t height, char* ws,
                        const Workspace& L, int32_t* status, cudaStream_t s, SortedPlan* out) {
  const int passes = sort_passes(height);
  // small batches: the whole sort in one cluster launch (cluster_sort.cu)
  static const bool force_global = std::getenv("B200EMB_FORCE_GLOBAL_SORT") != nullptr;
  if (!force_global && cluster_sort_fits(n)) {
    cudaError_t ce = launch_cluster_sort(
        rows, n, height, passes, reinterpret_cast<uint32_t*>(ws + L.off_keys[passes & 1]),
        reinterpret_cast<uint32_t*>(ws + L.off_perm[passes & 1]), status, s);
    if (ce == cudaSuccess) {
      if (out) *out = plan_view(height, ws, L);
      return cudaSuccess;
    }
    if (ce != cudaErrorNotSupported) return ce;
  }
  cudaError_t e = cudaMemsetAsync(ws, 0, L.zero_bytes, s);
  if (e != cudaSuccess) return e;
  uint32_t* keys[2] = {reinterpret_cast<uint32_t*>(ws + L.off_keys[0]),
                       reinterpret_cast<uint32_t*>(ws + L.off_keys[1])};
  uint32_t* perm[2] = {reinterpret_cast<uint32_t*>(ws + L.off_perm[0]),
                       reinterpret_cast<uint32_t*>(ws + L.off_perm[1])};
  uint32_t* ghist = reinterpret_cast<uint32_t*>(ws + L.off_hist);
  uint32_t* tickets = reinterpret_cast<uint32_t*>(ws + L.off_tickets);
  uint32_t* lookback = reinterpret_cast<uint32_t*>(ws + L.off_lookback);
  const int ni = static_cast<int>(n);
  const int prep_grid = static_cast<int>(std::min<int64_t>(ceil_div(n, 256 * 4), 148 * 4));
  k_prepare<<<prep_grid, 256, 0, s>>>(rows, ni, static_cast<uint32_t>(height), passes, keys[0],
                                      ghist, status);
  const int tiles = static_cast<int>(L.sort_tiles);
  for (int p = 0; p < passes; ++p) {
    uint32_t* st = lookback + static_cast<size_t>(p) * L.sort_tiles * kRadix;
    if (p == 0)
      k_radix_pass<false><<<tiles, kSortThreads, 0, s>>>(keys[0], nullptr, keys[1], perm[1], ni, 0,
                                                         ghist, tickets + p, st);
    else
      k_radix_pass<true><<<tiles, kSortThreads, 0, s>>>(keys[p & 1], perm[p & 1], keys[(p + 1) & 1],
                                                        perm[(p + 1) & 1], ni, p * kRadixBits,
                                                        ghist + p * kRadix, tickets + p, st);
  }
  if (out) *out = plan_view(height, ws, L);
  return cudaGetLastError();
}

cudaError_t launch_head_scan(const SortedPlan& plan, int64_t n, uint32_t invalid_key, char* ws,
                             const Workspace& L, int64_t* out_rows, int64_t* n_unique_out,
                             cudaStream_t s) {
  uint32_t* tickets = reinterpret_cast<uint32_t*>(ws + L.off_tickets);
  uint32_t* state = reinterpret_cast<uint32_t*>(ws + L.off_scan_state);
  // the scan may run more than once per plan (merge_add after plan_rows): reset its state
  cudaError_t e = cudaMemsetAsync(state, 0, sizeof(uint32_t) * L.scan_tiles, s);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(tickets + kMaxPasses, 0, sizeof(uint32_t), s);
  if (e != cudaSuccess) return e;
  k_head_scan<<<static_cast<int>(L.scan_tiles), 256, 0, s>>>(
      plan.keys, static_cast<int>(n), invalid_key, tickets + kMaxPasses, state,
      reinterpret_cast<uint32_t*>(ws + L.off_uidx), reinterpret_cast<uint32_t*>(ws + L.off_seg_start),
      out_rows, reinterpret_cast<PlanHeader*>(ws + L.off_plan), n_unique_out,
      static_cast<int>(L.scan_tiles));
  return cudaGetLastError();
}

}  // namespace b200emb
