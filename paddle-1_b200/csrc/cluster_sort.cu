// cluster_sort.cu — single-launch sort plan for small batches (N <= ~139k ids), sm_100a.
//
// At BASELINE's batch sizes (4096 x 26 = 106 496 ids per GPU) a whole step is ~15 us of HBM time,
// so a multi-launch global-memory radix sort (one launch + one chained scan per digit) costs more
// than everything else together.  Here ONE thread-block cluster of 16 CTAs keeps the whole
// (key, position) array in its distributed shared memory (16 x ~110 KB) and runs every LSD pass
// inside the cluster: per-warp match ranking (stable), per-CTA digit histograms exchanged by
// DSMEM loads, scatter by DSMEM stores, cluster.sync() instead of kernel boundaries.  Global
// memory is touched twice: ids in, sorted keys + positions out.  The other 132 SMs stay free for
// the forward gather running concurrently on another stream.
//
// Same contract as sort.cu's launch_plan: stable, out-of-range ids become key `height` and set
// status bit 1, results land in the workspace buffers plan_view() points at.
#include <cooperative_groups.h>

#include "common.cuh"
#include "internal.h"

namespace cg = cooperative_groups;

namespace b200emb {
namespace {

constexpr int kCsCluster = 16;
constexpr int kCsThreads = 512;
constexpr int kCsWarps = kCsThreads / 32;
constexpr int kCsArrays = 6;         // key/perm ping-pong (4) + key/perm staging (2)
constexpr int kCsMaxPerCta = 8704;   // 6 u32 arrays x 8704 x 4 B = 204 KB of the 227 KB
constexpr int kCsAuxWords = kCsWarps * kRadix + 3 * kRadix + 32;

__global__ void __launch_bounds__(kCsThreads, 1)
k_cluster_sort(const int64_t* __restrict__ rows, int n, uint32_t height, int passes, int m_cap,
               uint32_t* __restrict__ keys_out, uint32_t* __restrict__ perm_out, int32_t* status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // [key buf0 | key buf1 | perm buf0 | perm buf1 | key stage | perm stage], m_cap words each;
  // indexed arithmetically so the accesses stay LDS/STS rather than generic LD/ST
  uint32_t* const s_arr = reinterpret_cast<uint32_t*>(smem_raw);
  uint32_t* const s_stage_key = s_arr + 4 * m_cap;
  uint32_t* const s_stage_perm = s_arr + 5 * m_cap;
  uint32_t* s_wcnt = s_arr + kCsArrays * m_cap;     // [kCsWarps][kRadix] running counters
  uint32_t* s_hist = s_wcnt + kCsWarps * kRadix;    // [kRadix] this CTA's digit totals
  uint32_t* s_lstart = s_hist + kRadix;             // [kRadix] first local sorted slot of a digit
  uint32_t* s_gstart = s_lstart + kRadix;           // [kRadix] global position of that slot
  uint32_t* s_scan = s_gstart + kRadix;             // [8+8] warp sums of the two bin scans

  cg::cluster_group cluster = cg::this_cluster();
  const int rank = static_cast<int>(cluster.block_rank());
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int start = rank * m_cap;
  const int m = max(0, min(m_cap, n - start));

  bool bad = false;
  for (int i = tid; i < m; i += kCsThreads) {
    const int64_t r = rows[start + i];
    const bool ok = r >= 0 && r < static_cast<int64_t>(height);
    bad |= !ok;
    s_arr[i] = ok ? static_cast<uint32_t>(r) : height;
    s_arr[2 * m_cap + i] = static_cast<uint32_t>(start + i);
  }
  if (bad && status) atomicOr(status, 2);

  const int per_warp = ((m_cap + kCsWarps - 1) / kCsWarps + 31) / 32 * 32;
  const int wbase = warp * per_warp;
  const int rounds = per_warp / 32;
  const uint32_t lt = lanemask_lt();
  int cur = 0;

  for (int p = 0; p < passes; ++p) {
    const int shift = p * kRadixBits;
    const uint32_t* src_key = s_arr + cur * m_cap;
    const uint32_t* src_perm = s_arr + (2 + cur) * m_cap;
    for (int i = tid; i < kCsWarps * kRadix; i += kCsThreads) s_wcnt[i] = 0;
    __syncthreads();
    // phase A: per-warp digit counts
    for (int j = 0; j < rounds; ++j) {
      const int idx = wbase + j * 32 + lane;
      const bool valid = idx < m;
      const uint32_t digit = valid ? ((src_key[idx] >> shift) & (kRadix - 1)) : 0u;
      const uint32_t peers = match_digit<kRadixBits>(digit, valid);
      if (valid && lane == __ffs(peers) - 1) s_wcnt[warp * kRadix + digit] += __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    uint32_t tot = 0, lincl = 0;
    if (tid < kRadix) {
#pragma unroll
      for (int w = 0; w < kCsWarps; ++w) {
        const uint32_t c = s_wcnt[w * kRadix + tid];
        s_wcnt[w * kRadix + tid] = tot;
        tot += c;
      }
      s_hist[tid] = tot;
      lincl = tot;   // local exclusive scan over bins -> slot of the digit's first element
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, lincl, o);
        if (lane >= o) lincl += up;
      }
      if (lane == 31) s_scan[warp] = lincl;
    }
    cluster.barrier_arrive();   // s_hist published; overlap the local scatter with the barrier
    __syncthreads();
    if (tid < kRadix) {
      uint32_t off = 0;
#pragma unroll
      for (int w = 0; w < kRadix / 32; ++w) off += (w < warp) ? s_scan[w] : 0u;
      const uint32_t ls = off + lincl - tot;
      s_lstart[tid] = ls;
#pragma unroll
      for (int w = 0; w < kCsWarps; ++w) s_wcnt[w * kRadix + tid] += ls;
    }
    __syncthreads();
    // phase B: stable LOCAL scatter into the staging buffers (sorted by digit inside the CTA)
    for (int j = 0; j < rounds; ++j) {
      const int idx = wbase + j * 32 + lane;
      const bool valid = idx < m;
      const uint32_t key = valid ? src_key[idx] : 0u;
      const uint32_t pm = valid ? src_perm[idx] : 0u;
      const uint32_t digit = valid ? ((key >> shift) & (kRadix - 1)) : 0u;
      const uint32_t peers = match_digit<kRadixBits>(digit, valid);
      const int leader = valid ? __ffs(peers) - 1 : lane;
      uint32_t old = 0;
      if (valid && lane == leader) {
        old = s_wcnt[warp * kRadix + digit];
        s_wcnt[warp * kRadix + digit] = old + __popc(peers);
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      if (valid) {
        const uint32_t lpos = old + __popc(peers & lt);
        s_stage_key[lpos] = key;
        s_stage_perm[lpos] = pm;
      }
      __syncwarp();
    }
    cluster.barrier_wait();     // every CTA's s_hist is now readable
    uint32_t total = 0, before = 0, gincl = 0;
    if (tid < kRadix) {
#pragma unroll
      for (int c = 0; c < kCsCluster; ++c) {
        const uint32_t h = *cluster.map_shared_rank(&s_hist[tid], c);
        total += h;
        before += (c < rank) ? h : 0u;
      }
      gincl = total;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, gincl, o);
        if (lane >= o) gincl += up;
      }
      if (lane == 31) s_scan[8 + warp] = gincl;
    }
    __syncthreads();   // also orders phase B's staging writes before phase C's reads
    if (tid < kRadix) {
      uint32_t off = 0;
#pragma unroll
      for (int w = 0; w < kRadix / 32; ++w) off += (w < warp) ? s_scan[8 + w] : 0u;
      s_gstart[tid] = off + gincl - total + before;
    }
    __syncthreads();
    // phase C: copy the digit runs out to their owners; consecutive lanes -> consecutive words
    uint32_t* dst_key = s_arr + (cur ^ 1) * m_cap;
    uint32_t* dst_perm = s_arr + (2 + (cur ^ 1)) * m_cap;
    for (int i = tid; i < m; i += kCsThreads) {
      const uint32_t key = s_stage_key[i];
      const uint32_t pm = s_stage_perm[i];
      const uint32_t digit = (key >> shift) & (kRadix - 1);
      const uint32_t pos = s_gstart[digit] + (static_cast<uint32_t>(i) - s_lstart[digit]);
      const uint32_t r = pos / static_cast<uint32_t>(m_cap);
      const uint32_t off = pos - r * static_cast<uint32_t>(m_cap);
      cluster.map_shared_rank(dst_key, r)[off] = key;
      cluster.map_shared_rank(dst_perm, r)[off] = pm;
    }
    cluster.sync();
    cur ^= 1;
  }
  for (int i = tid; i < m; i += kCsThreads) {
    keys_out[start + i] = s_arr[cur * m_cap + i];
    perm_out[start + i] = s_arr[(2 + cur) * m_cap + i];
  }
}

}  // namespace

bool cluster_sort_fits(int64_t n) {
  return n > 0 && ceil_div(n, kCsCluster) <= kCsMaxPerCta;
}

// Returns cudaErrorNotSupported when this device/driver cannot co-schedule the 16-CTA cluster,
// in which case the caller falls back to the multi-launch global sort.
cudaError_t launch_cluster_sort(const int64_t* rows, int64_t n, int64_t height, int passes,
                                uint32_t* keys_out, uint32_t* perm_out, int32_t* status,
                                cudaStream_t s) {
  const int m_cap = static_cast<int>((ceil_div(n, kCsCluster) + 31) / 32 * 32);
  const size_t smem = sizeof(uint32_t) * (kCsArrays * static_cast<size_t>(m_cap) + kCsAuxWords);
  cudaError_t e = cudaFuncSetAttribute(k_cluster_sort, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(smem));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_cluster_sort, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kCsCluster);
  cfg.blockDim = dim3(kCsThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCsCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static thread_local int supported[64] = {};  // per device: 0 unknown, 1 yes, -1 no
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && supported[dev] == 0) {
    int clusters = 0;
    cudaLaunchConfig_t probe = cfg;
    probe.dynamicSmemBytes = sizeof(uint32_t) * (kCsArrays * static_cast<size_t>(kCsMaxPerCta) + kCsAuxWords);
    cudaFuncSetAttribute(k_cluster_sort, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(probe.dynamicSmemBytes));
    e = cudaOccupancyMaxActiveClusters(&clusters, k_cluster_sort, &probe);
    supported[dev] = (e == cudaSuccess && clusters >= 1) ? 1 : -1;
    if (e != cudaSuccess) cudaGetLastError();
  }
  if (dev < 0 || dev >= 64 || supported[dev] < 0) return cudaErrorNotSupported;
  const int ni = static_cast<int>(n);
  const uint32_t h = static_cast<uint32_t>(height);
  return cudaLaunchKernelEx(&cfg, k_cluster_sort, rows, ni, h, passes, m_cap, keys_out, perm_out,
                            status);
}

}  // namespace b200emb
