// sort.cu — deterministic sort-by-id ("plan") for SelectedRows rows, sm_100a.
//
// Replaces the host-side std::set + per-row linear scan of the reference's MergeAdd
// (operators/math/selected_rows_functor.cc:193-228, .cu:235-297) with a stable LSD radix
// sort of (id -> original position) done entirely on the device:
//   k_prepare_direct   int64 ids -> u32 keys (+ range check) and the first pass's per-tile digit
//                      counts H[0][tile][bin]; clears the later passes' matrices;
//   k_colscan          (above 64 tiles) turns H[pass] into exclusive prefixes over tiles;
//   k_radix_pass_direct  one launch per digit of up to 10 bits: warp-level match ranking
//                      (stable), scatter, and — while scattering — the NEXT pass's H counted with
//                      integer reductions keyed by (destination tile, next digit);
//   k_head_count / k_head_write   run-length heads of the sorted keys -> unique index of every
//                      element, first position of every unique row, ascending unique rows
//                      (= std::set order), U.
// Only ceil(log2(height+1)) key bits are sorted.  The kernel boundary is the only
// synchronisation: no flags, no polling, no look-back, no memset (an earlier chained-scan
// version lost 12-17k cycles per tile to look-back waits at 832 tiles; profiles/r1_notes.md).
// The sort moves 8 B per id per pass, which is noise next to the 4*D B of row data each id
// drags through HBM; launch count is what matters for small batches, scatter throughput
// (~40 G key-passes/s) for large ones.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace b200emb {

namespace {
// plain (non-polling) vector load of K counters, L2-cached only
template <int K>
__device__ __forceinline__ void ld_cg_vec(const uint32_t* p, uint32_t (&v)[K]) {
  if constexpr (K == 1) {
    v[0] = __ldcg(p);
  } else if constexpr (K == 2) {
    const uint2 t = __ldcg(reinterpret_cast<const uint2*>(p));
    v[0] = t.x; v[1] = t.y;
  } else {
#pragma unroll
    for (int q = 0; q < K / 4; ++q) {
      const uint4 t = __ldcg(reinterpret_cast<const uint4*>(p) + q);
      v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
    }
  }
}

// ---- small batches (<= kDirectMaxTiles tiles): no chained scan at all --------------------
// With at most 64 tiles every tile's per-digit counts fit a small matrix H[pass][tile][bin] in
// L2.  k_prepare_direct fills H[0] (one CTA per tile, plain stores) and clears the later
// passes' matrices; each pass kernel reads the whole column block of its bins (no flags, no
// polling: the kernel boundary is the synchronisation), which yields both the exclusive prefix
// over earlier tiles and the global digit totals, and — while scattering — counts the NEXT pass's
// matrix with fire-and-forget integer reductions keyed by (destination tile, next digit).
// Launches per plan: 1 + passes, no memset, no look-back.
constexpr int kDirectMaxTiles = 64;

template <int BITS, int ITEMS>
__global__ void __launch_bounds__(kSortThreads, 4)
k_prepare_direct(const int64_t* __restrict__ rows, int n, uint32_t height, KeyMap km, int passes,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ hmat, size_t per_pass,
                 uint32_t* __restrict__ arrive, int arrive_n, int32_t* status) {
  constexpr int kBins = 1 << BITS;
  __shared__ uint32_t sh[kBins];
  const int tid = threadIdx.x, lane = tid & 31;
  const int tile = blockIdx.x;
  // the chunked reduce's chain arrival counters (segreduce.cu) start every plan at zero
  for (int i = tile * kSortThreads + tid; i < arrive_n; i += gridDim.x * kSortThreads) arrive[i] = 0;
  for (int i = tid; i < kBins; i += kSortThreads) sh[i] = 0;
  __syncthreads();
  bool bad = false;
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const int i = tile * (ITEMS * kSortThreads) + j * kSortThreads + tid;
    const bool valid = i < n;
    uint32_t key = 0;
    if (valid) {
      const int64_t r = rows[i];
      const bool ok = r >= 0 && r < static_cast<int64_t>(km.raw_height);
      key = ok ? km(static_cast<uint32_t>(r)) : height;
      bad |= !ok;
      keys[i] = key;
    }
    const uint32_t digit = key & (kBins - 1);
    const uint32_t peers = match_digit<BITS>(digit, valid);
    if (valid && lane == __ffs(peers) - 1) atomicAdd(&sh[digit], __popc(peers));
  }
  __syncthreads();
  for (int i = tid; i < kBins; i += kSortThreads) {
    hmat[static_cast<size_t>(tile) * kBins + i] = sh[i];
    for (int p = 1; p < passes; ++p)
      hmat[p * per_pass + static_cast<size_t>(tile) * kBins + i] = 0;
  }
  if (bad && status) atomicOr(status, 2);
}

// ---- larger batches: the same matrix, scanned by its own small launch -----------------------
// Reading a whole column block per tile is quadratic in the tile count, and a chained look-back
// makes every resident tile poll its predecessors' counters (measured: 12-17k cycles of waiting
// per tile at 832 tiles).  Instead k_colscan turns H[pass] into exclusive prefixes over tiles in
// place — one CTA per 8 bins, a warp instruction covers 4 tiles x 8 bins (32-byte sectors),
// 32 warps split the tiles — and leaves the column totals in row `tiles`.  The pass kernel then
// reads exactly two rows.  Launches per plan: 1 + 2*passes, no flags, no polling, no memset.
constexpr int kColscanBins = 8;
constexpr int kColscanThreads = 1024;

__global__ void __launch_bounds__(kColscanThreads)
k_colscan(uint32_t* __restrict__ hmat, int tiles, int bins) {
  constexpr int kWarps = kColscanThreads / 32;
  constexpr int kU = 4;                               // row groups in flight per thread
  __shared__ uint32_t s_tot[kWarps][kColscanBins];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = lane & 7, r = lane >> 3;
  const int groups = (tiles + 3) / 4;                  // groups of 4 consecutive tiles
  const int per_warp = (groups + kWarps - 1) / kWarps;
  const int g_lo = warp * per_warp, g_hi = min(groups, g_lo + per_warp);
  uint32_t* col = hmat + blockIdx.x * kColscanBins + b;
  uint32_t sum = 0;
  for (int g = g_lo; g < g_hi; g += kU) {
    uint32_t v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int t = (g + u) * 4 + r;
      v[u] = (g + u < g_hi && t < tiles) ? col[static_cast<size_t>(t) * bins] : 0u;
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) sum += v[u];
  }
  sum += __shfl_xor_sync(0xffffffffu, sum, 8);
  sum += __shfl_xor_sync(0xffffffffu, sum, 16);
  if (r == 0) s_tot[warp][b] = sum;
  __syncthreads();
  uint32_t run = 0, total = 0;
#pragma unroll 8
  for (int w = 0; w < kWarps; ++w) {
    const uint32_t c = s_tot[w][b];
    run += w < warp ? c : 0u;
    total += c;
  }
  for (int g = g_lo; g < g_hi; g += kU) {
    uint32_t v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int t = (g + u) * 4 + r;
      v[u] = (g + u < g_hi && t < tiles) ? col[static_cast<size_t>(t) * bins] : 0u;
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const uint32_t v0 = __shfl_sync(0xffffffffu, v[u], b);
      const uint32_t v1 = __shfl_sync(0xffffffffu, v[u], 8 + b);
      const uint32_t v2 = __shfl_sync(0xffffffffu, v[u], 16 + b);
      const uint32_t v3 = __shfl_sync(0xffffffffu, v[u], 24 + b);
      const uint32_t pre = (r > 0 ? v0 : 0u) + (r > 1 ? v1 : 0u) + (r > 2 ? v2 : 0u);
      const int t = (g + u) * 4 + r;
      if (g + u < g_hi && t < tiles) col[static_cast<size_t>(t) * bins] = run + pre;
      run += v0 + v1 + v2 + v3;
    }
  }
  if (warp == 0 && r == 0) col[static_cast<size_t>(tiles) * bins] = total;
}

template <int BITS, bool HAS_PERM_IN, bool SCANNED, int ITEMS>
__global__ void __launch_bounds__(kSortThreads, 4)   // <= 64 registers: fits beside the row kernels on an SM
k_radix_pass_direct(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ perm_in,
                    uint32_t* __restrict__ keys_out, uint32_t* __restrict__ perm_out, int n,
                    int shift, const uint32_t* __restrict__ hmat, uint32_t* __restrict__ hnext) {
  constexpr int kWarps = kSortThreads / 32;
  constexpr int kBins = 1 << BITS;
  constexpr int K = kBins / kSortThreads;
  __shared__ uint32_t s_base[kBins];
  __shared__ uint16_t s_cnt[kWarps * kBins];
  __shared__ uint32_t s_warp_sum[kWarps];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x, tiles = gridDim.x;
  {
    uint32_t* z = reinterpret_cast<uint32_t*>(s_cnt);
    for (int i = tid; i < kWarps * kBins / 2; i += kSortThreads) z[i] = 0;
  }
  const int warp_start = tile * (ITEMS * kSortThreads) + warp * (32 * ITEMS);
  uint16_t* my_cnt = s_cnt + warp * kBins;
  uint32_t key[ITEMS], pm[ITEMS], rank[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const int idx = warp_start + j * 32 + lane;
    const bool valid = idx < n;
    key[j] = valid ? keys_in[idx] : 0u;
    pm[j] = valid ? (HAS_PERM_IN ? perm_in[idx] : static_cast<uint32_t>(idx)) : 0u;
  }
  // this thread's K bins: prefix over earlier tiles + totals over all tiles
  uint32_t excl[K], total[K];
  if constexpr (SCANNED) {   // k_colscan left the prefixes in row `tile`, the totals in row `tiles`
    ld_cg_vec<K>(hmat + static_cast<size_t>(tile) * kBins + tid * K, excl);
    ld_cg_vec<K>(hmat + static_cast<size_t>(tiles) * kBins + tid * K, total);
  } else {
#pragma unroll
    for (int k = 0; k < K; ++k) { excl[k] = 0; total[k] = 0; }
    constexpr int kU = K >= 4 ? 26 : 32;   // rows of the matrix in flight per round trip
    for (int t0 = 0; t0 < tiles; t0 += kU) {
      uint32_t v[kU][K];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int t = t0 + u;
        if (t < tiles) {
          ld_cg_vec<K>(hmat + static_cast<size_t>(t) * kBins + tid * K, v[u]);
        } else {
#pragma unroll
          for (int k = 0; k < K; ++k) v[u][k] = 0;
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u)
#pragma unroll
        for (int k = 0; k < K; ++k) {
          total[k] += v[u][k];
          excl[k] += (t0 + u < tile) ? v[u][k] : 0u;
        }
    }
  }
  __syncthreads();
  // stable rank inside the warp: items are ordered (j, lane) == ascending idx
  const uint32_t lt = lanemask_lt();
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const bool valid = (warp_start + j * 32 + lane) < n;
    const uint32_t digit = (key[j] >> shift) & (kBins - 1);
    const uint32_t peers = match_digit<BITS>(digit, valid);
    const int leader = valid ? __ffs(peers) - 1 : lane;
    uint32_t old = 0;
    if (valid && lane == leader) {
      old = my_cnt[digit];
      my_cnt[digit] = static_cast<uint16_t>(old + __popc(peers));
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[j] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  // exclusive offsets of each warp inside the tile
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int bin = tid * K + k;
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const uint32_t c = s_cnt[w * kBins + bin];
      s_cnt[w * kBins + bin] = static_cast<uint16_t>(t);
      t += c;
    }
  }
  // exclusive scan of the digit totals over bins
  uint32_t gsum = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) gsum += total[k];
  uint32_t incl = gsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_warp_sum[warp] = incl;
  __syncthreads();
  uint32_t run = incl - gsum;
#pragma unroll
  for (int w = 0; w < kWarps; ++w) run += (w < warp) ? s_warp_sum[w] : 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    s_base[tid * K + k] = run + excl[k];
    run += total[k];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const int idx = warp_start + j * 32 + lane;
    if (idx < n) {
      const uint32_t digit = (key[j] >> shift) & (kBins - 1);
      const uint32_t pos = s_base[digit] + my_cnt[digit] + rank[j];
      keys_out[pos] = key[j];
      perm_out[pos] = pm[j];
      if (hnext) {
        const uint32_t nd = (key[j] >> (shift + BITS)) & (kBins - 1);
        atomicAdd(hnext + static_cast<size_t>(pos / (ITEMS * kSortThreads)) * kBins + nd, 1u);
      }
    }
  }
}

// Run-length heads of the sorted keys, in two launches with the kernel boundary as the only
// synchronisation: k_head_count leaves every tile's head count in counts[], k_head_write sums
// the counts of the tiles before its own (a coalesced read of <= tiles words; tiles <= a few
// thousand for the batches of BASELINE.json) and writes uidx / seg_start / unique rows / U.
__device__ __forceinline__ uint32_t tile_heads(const uint32_t* __restrict__ skey, int n,
                                               uint32_t invalid_key, int idx0, uint32_t (&k)[kScanTile / 256],
                                               uint32_t& headmask) {
  constexpr int kItems = kScanTile / 256;
  static_assert(kItems == 8, "two 16-byte loads per thread");
  uint32_t prev = (idx0 > 0 && idx0 - 1 < n) ? skey[idx0 - 1] : kNoKey;
  if (idx0 + kItems <= n) {
    const uint4 a = *reinterpret_cast<const uint4*>(skey + idx0);
    const uint4 b = *reinterpret_cast<const uint4*>(skey + idx0 + 4);
    k[0] = a.x; k[1] = a.y; k[2] = a.z; k[3] = a.w; k[4] = b.x; k[5] = b.y; k[6] = b.z; k[7] = b.w;
  } else {
#pragma unroll
    for (int j = 0; j < kItems; ++j) k[j] = idx0 + j < n ? skey[idx0 + j] : kNoKey;
  }
  uint32_t cnt = 0;
  headmask = 0;
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const bool head = idx0 + j < n && k[j] != prev && k[j] != invalid_key;
    headmask |= head ? (1u << j) : 0u;
    cnt += head;
    prev = k[j];
  }
  return cnt;
}

__global__ void __launch_bounds__(256)
k_head_count(const uint32_t* __restrict__ skey, int n, uint32_t invalid_key,
             uint32_t* __restrict__ counts) {
  uint32_t k[kScanTile / 256], headmask;
  const int idx0 = blockIdx.x * kScanTile + threadIdx.x * (kScanTile / 256);
  const uint32_t cnt = tile_heads(skey, n, invalid_key, idx0, k, headmask);
  const uint32_t w = __reduce_add_sync(0xffffffffu, cnt);
  __shared__ uint32_t s_w[8];
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = w;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s_w[i];
    counts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256)
k_head_write(const uint32_t* __restrict__ skey, int n, uint32_t invalid_key,
             const uint32_t* __restrict__ counts, uint32_t* __restrict__ uidx,
             uint32_t* __restrict__ seg_start, int64_t* __restrict__ out_rows, PlanHeader* hdr,
             int64_t* n_unique_out) {
  constexpr int kItems = kScanTile / 256;
  __shared__ uint32_t s_warp_sum[8], s_warp_pre[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  uint32_t before = 0;                       // heads in earlier tiles
  for (int t = tid; t < tile; t += 256) before += counts[t];
  uint32_t k[kItems], headmask;
  const int idx0 = tile * kScanTile + tid * kItems;
  const uint32_t cnt = tile_heads(skey, n, invalid_key, idx0, k, headmask);
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  before = __reduce_add_sync(0xffffffffu, before);
  if (lane == 31) s_warp_sum[warp] = incl;
  if (lane == 0) s_warp_pre[warp] = before;
  __syncthreads();
  uint32_t warp_off = 0, block_total = 0, tile_excl = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    warp_off += (w < warp) ? s_warp_sum[w] : 0u;
    block_total += s_warp_sum[w];
    tile_excl += s_warp_pre[w];
  }
  if (tid == 0 && tile == static_cast<int>(gridDim.x) - 1) {
    hdr->n_unique = tile_excl + block_total;
    if (n_unique_out) *n_unique_out = static_cast<int64_t>(tile_excl + block_total);
  }
  uint32_t run = tile_excl + warp_off + incl - cnt;  // heads before this thread's items
  uint32_t u[kItems];
#pragma unroll
  for (int j = 0; j < kItems; ++j) {
    const bool head = (headmask >> j) & 1u;
    run += head;
    // out-of-range ids (key == invalid_key) belong to no unique row: kNoKey tells expand_unique / the
    // inverse map to hand back a zero row instead of some other id's embedding
    u[j] = k[j] == invalid_key ? kNoKey : run - 1u;
    if (head && idx0 + j < n) {
      seg_start[u[j]] = static_cast<uint32_t>(idx0 + j);
      if (out_rows) out_rows[u[j]] = static_cast<int64_t>(k[j]);
    }
  }
  if (idx0 + kItems <= n) {
    *reinterpret_cast<uint4*>(uidx + idx0) = make_uint4(u[0], u[1], u[2], u[3]);
    *reinterpret_cast<uint4*>(uidx + idx0 + 4) = make_uint4(u[4], u[5], u[6], u[7]);
  } else {
#pragma unroll
    for (int j = 0; j < kItems; ++j)
      if (idx0 + j < n) uidx[idx0 + j] = u[j];
  }
}

int key_bits(int64_t height) {  // keys span [0, height] (height = the invalid sentinel)
  int b = 1;
  while ((1ull << b) <= static_cast<unsigned long long>(height)) ++b;
  return b;
}

// Digits of up to 10 bits (20 KB of shared counters per CTA) to save passes; the fewest passes
// that cover the key, then the narrowest digit that still fits them.
constexpr int kMaxDigitBits = 10;
// Keys per tile.  Batches between the direct mode and about two resident waves of 2048-key tiles are
// latency-bound with half-empty SMs (320 tiles at C3: 15 % issue utilisation): 1024-key tiles double
// the CTAs in flight there.  Larger batches fill the chip anyway and keep the smaller count matrix.
constexpr int64_t kHalfTileMaxKeys = 148LL * 4 * 2 * 1024;
int64_t sort_tile_keys(int64_t n) {
  if (ceil_div(n, kSortTile) <= kDirectMaxTiles || n > kHalfTileMaxKeys) return kSortTile;
  return kSortTile / 2;
}
struct DigitPlan { int passes, bits; bool direct; };
DigitPlan digit_plan(int64_t n, int64_t height) {
  const int kb = key_bits(height);
  DigitPlan p;
  p.direct = ceil_div(n, kSortTile) <= kDirectMaxTiles;
  p.passes = static_cast<int>(ceil_div(kb, kMaxDigitBits));
  p.bits = static_cast<int>(ceil_div(kb, p.passes));
  if (p.bits < kRadixBits) p.bits = kRadixBits;
  return p;
}
// counter words per tile the workspace must hold for any height: passes x bins <= 3 x 1024
constexpr int64_t kDigitWordsPerTile = 3LL * (1 << kMaxDigitBits);
static_assert(kMaxPasses * (1 << kRadixBits) <= kDigitWordsPerTile, "4 passes of 8 bits");

}  // namespace

Workspace make_workspace(int64_t n, int64_t d, int64_t elem_size) {
  Workspace L{};
  if (n < 1) n = 1;
  L.sort_tiles = ceil_div(n, sort_tile_keys(n));
  L.scan_tiles = ceil_div(n, kScanTile);
  L.max_ctas = ceil_div(n, static_cast<int64_t>(kMinGroupsPerCta) * kMinChunk) + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  L.off_digit_counts = take(sizeof(uint32_t) * (L.sort_tiles + 1) * kDigitWordsPerTile);
  L.off_scan_state = take(sizeof(uint32_t) * L.scan_tiles);
  L.off_plan = take(sizeof(PlanHeader));
  for (int b = 0; b < 2; ++b) L.off_keys[b] = take(sizeof(uint32_t) * n);
  for (int b = 0; b < 2; ++b) L.off_perm[b] = take(sizeof(uint32_t) * n);
  L.off_uidx = take(sizeof(uint32_t) * n);
  L.off_seg_start = take(sizeof(uint32_t) * (n + 1));
  L.off_cta_arrive = take(sizeof(uint32_t) * L.max_ctas);
  L.off_cta_head = take(static_cast<size_t>(elem_size) * L.max_ctas * d);
  L.off_cta_tail = take(static_cast<size_t>(elem_size) * L.max_ctas * d);
  L.total = off;
  return L;
}

SortedPlan plan_view(int64_t n, int64_t height, char* ws, const Workspace& L) {
  SortedPlan p;
  p.passes = digit_plan(n, height).passes;
  p.keys = reinterpret_cast<const uint32_t*>(ws + L.off_keys[p.passes & 1]);
  p.perm = reinterpret_cast<const uint32_t*>(ws + L.off_perm[p.passes & 1]);
  return p;
}

namespace {
template <int BITS, bool SCANNED, int ITEMS>
cudaError_t launch_passes_direct(const int64_t* rows, int n, uint32_t height, KeyMap km, int passes,
                                 char* ws, const Workspace& L, int32_t* status, cudaStream_t s) {
  constexpr int kBins = 1 << BITS;
  uint32_t* keys[2] = {reinterpret_cast<uint32_t*>(ws + L.off_keys[0]),
                       reinterpret_cast<uint32_t*>(ws + L.off_keys[1])};
  uint32_t* perm[2] = {reinterpret_cast<uint32_t*>(ws + L.off_perm[0]),
                       reinterpret_cast<uint32_t*>(ws + L.off_perm[1])};
  uint32_t* hmat = reinterpret_cast<uint32_t*>(ws + L.off_digit_counts);
  const int tiles = static_cast<int>(L.sort_tiles);
  const size_t per_pass = static_cast<size_t>(tiles + 1) * kBins;   // + the totals row
  k_prepare_direct<BITS, ITEMS><<<tiles, kSortThreads, 0, s>>>(
      rows, n, height, km, passes, keys[0], hmat, per_pass,
      reinterpret_cast<uint32_t*>(ws + L.off_cta_arrive), static_cast<int>(L.max_ctas), status);
  for (int p = 0; p < passes; ++p) {
    uint32_t* hn = p + 1 < passes ? hmat + (p + 1) * per_pass : nullptr;
    if constexpr (SCANNED)
      k_colscan<<<kBins / kColscanBins, kColscanThreads, 0, s>>>(hmat + p * per_pass, tiles, kBins);
    if (p == 0)
      k_radix_pass_direct<BITS, false, SCANNED, ITEMS><<<tiles, kSortThreads, 0, s>>>(
          keys[0], nullptr, keys[1], perm[1], n, 0, hmat, hn);
    else
      k_radix_pass_direct<BITS, true, SCANNED, ITEMS><<<tiles, kSortThreads, 0, s>>>(
          keys[p & 1], perm[p & 1], keys[(p + 1) & 1], perm[(p + 1) & 1], n, p * BITS,
          hmat + p * per_pass, hn);
  }
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_plan(const int64_t* rows, int64_t n, int64_t height, char* ws,
                        const Workspace& L, int32_t* status, cudaStream_t s, SortedPlan* out,
                        const KeyMap* keymap) {
  KeyMap km;
  km.raw_height = static_cast<uint32_t>(height);
  if (keymap) km = *keymap;
  const DigitPlan dp = digit_plan(n, height);
  // optional: the whole sort in one 16-CTA cluster launch (cluster_sort.cu).  Measured slower
  // than the multi-launch path on B200, so it is opt-in.
  static const bool use_cluster = std::getenv("B200EMB_CLUSTER_SORT") != nullptr;
  if (use_cluster && keymap == nullptr && cluster_sort_fits(n)) {   // the cluster kernel sorts raw ids: no key map
    const int cpasses = static_cast<int>(ceil_div(key_bits(height), kRadixBits));
    // the cluster kernel always uses 8-bit digits; it writes straight into the buffers the
    // digit plan's parity selects
    cudaError_t ce = launch_cluster_sort(
        rows, n, height, cpasses, reinterpret_cast<uint32_t*>(ws + L.off_keys[dp.passes & 1]),
        reinterpret_cast<uint32_t*>(ws + L.off_perm[dp.passes & 1]), status, s);
    if (ce == cudaSuccess)
      ce = cudaMemsetAsync(ws + L.off_cta_arrive, 0, sizeof(uint32_t) * L.max_ctas, s);
    if (ce == cudaSuccess) {
      if (out) *out = plan_view(n, height, ws, L);
      return cudaSuccess;
    }
    if (ce != cudaErrorNotSupported) return ce;
  }
  const int ni = static_cast<int>(n);
  const uint32_t h = static_cast<uint32_t>(height);
  cudaError_t e;
  const bool half = sort_tile_keys(n) != kSortTile;
  if (dp.direct) {
    switch (dp.bits) {
      case 8: e = launch_passes_direct<8, false, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      case 9: e = launch_passes_direct<9, false, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      default: e = launch_passes_direct<10, false, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
    }
  } else if (half) {
    switch (dp.bits) {
      case 8: e = launch_passes_direct<8, true, kSortItems / 2>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      case 9: e = launch_passes_direct<9, true, kSortItems / 2>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      default: e = launch_passes_direct<10, true, kSortItems / 2>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
    }
  } else {
    switch (dp.bits) {
      case 8: e = launch_passes_direct<8, true, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      case 9: e = launch_passes_direct<9, true, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
      default: e = launch_passes_direct<10, true, kSortItems>(rows, ni, h, km, dp.passes, ws, L, status, s); break;
    }
  }
  if (e != cudaSuccess) return e;
  if (out) *out = plan_view(n, height, ws, L);
  return cudaSuccess;
}

cudaError_t launch_head_scan(const SortedPlan& plan, int64_t n, uint32_t invalid_key, char* ws,
                             const Workspace& L, int64_t* out_rows, int64_t* n_unique_out,
                             cudaStream_t s) {
  uint32_t* counts = reinterpret_cast<uint32_t*>(ws + L.off_scan_state);
  const int tiles = static_cast<int>(L.scan_tiles);
  k_head_count<<<tiles, 256, 0, s>>>(plan.keys, static_cast<int>(n), invalid_key, counts);
  k_head_write<<<tiles, 256, 0, s>>>(
      plan.keys, static_cast<int>(n), invalid_key, counts,
      reinterpret_cast<uint32_t*>(ws + L.off_uidx), reinterpret_cast<uint32_t*>(ws + L.off_seg_start),
      out_rows, reinterpret_cast<PlanHeader*>(ws + L.off_plan), n_unique_out);
  return cudaGetLastError();
}

}  // namespace b200emb
