// shard_engine.cu — the row-sharded embedding step behind the C ABI (b200emb_shard_*, include/b200emb.h).
//
// Replaces ParallelExecutor's multi-device SelectedRows path — ReduceOpHandle gathers every device's sparse
// gradient on one device (framework/details/reduce_op_handle.cc:88-94, reduce_and_gather.h:56-93), the
// optimizer runs there, BroadcastOpHandle sends the parameter back (broadcast_op_handle.cc:71-122) — with the
// table sharded by row, owner(id) = id % G, local row = id / G (operators/split_ids_op.h:45-49), results in
// the requester's order (operators/merge_ids_op.h:63-87).
//
// One engine per rank (one GPU).  Every rank owns one peer-visible block (CUDA IPC between processes, plain
// peer pointers inside one process) that the OTHER ranks' kernels store into over NVLink:
//     header    int64 [8]               {magic, G, capacities, D}: connect() refuses ranks whose layout differs
//     ids_in    int64 [4][G][cap_src]   per prepare slot: unique local rows requested by each source rank
//     meta_in   int64 [4][G][2]         per slot: {count, offset of these rows in the requester's unique list}
//     flags     u64   [4][3][G]         per slot: arrival counters of ids (by source), rows (by owner), gradients (by source)
//     meta_back int64 [G][2]            per owner: {where this rank's rows start in the owner's compact list, rows accepted}
//     uvals     T     [cap_ids][D]      unique rows returned by the owners, in this rank's unique order
//     grads_in  T     [cap_recv][D]     merged gradient rows of every source, compact, source-major
// A step is a fixed sequence of kernels per rank; nothing is read on the host and nothing waits for ALL
// ranks: a consumer kernel spins only on the flags of the ranks whose data it is about to read.  Flags carry
// the use count of the slot (seq[slot], advanced by prepare), so they never need clearing.
//     prepare  (ids only: runs a step ahead on a side stream)
//              sort plan + de-duplication + split points, then
//              k_send_ids        unique local rows -> owners' ids_in[slot], meta, ids flag   [NVLink 8 B / unique id]
//     lookup   k_owner_gather    (owner) wait ids flags; rows of the shard -> requesters' uvals; the received ids
//                                compacted (and padded) for the update's sort plan; meta_back + rows flags from
//                                the same kernel (every CTA fences its stores)              [NVLink 4D B / unique id]
//              sort plan of the received rows (side stream: only the update needs it)
//              k_permute_rows    wait rows flags; expand unique rows to batch order (= merge_ids)
//     update   k_chunk_reduce<MergeAdd>  local duplicate reduction -> owners' grads_in, then the gradient flags
//                                from the same kernel                                        [NVLink 4D B / unique id]
//              k_chunk_reduce<Sgd|Adam>  (owner) wait gradient flags; fused update of the shard
// The big buffers are single: the step's own data dependencies order their reuse (a rank cannot store the
// gradients of step i+1 before it has expanded the rows of step i+1, which every owner sent after its update of
// step i, ...); the id buffers are per slot because prepare() of step i+1 runs beside step i.
#include "../../include/b200emb.h"

#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "common.cuh"
#include "internal.h"

using namespace b200emb;

namespace {

constexpr int kMaxWorld = 64;
constexpr int kPrepSlots = 4;

struct BlockLayout {
  size_t off_header, off_ids_in, off_meta_in, off_meta_back, off_flags, off_uvals, off_grads_in, total;
  size_t ids_slot_bytes, meta_slot_bytes, flags_slot_bytes;   // strides between prepare slots
};
constexpr long long kBlockMagic = 0x4232303045534832LL;   // every rank's block must have the same layout

BlockLayout make_block(int world, int64_t cap_src, int64_t cap_ids, int64_t cap_recv, int64_t d) {
  BlockLayout B{};
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  B.off_header = take(sizeof(long long) * 8);   // {magic, world, cap_src, cap_ids, cap_recv, d}
  B.ids_slot_bytes = align_up(sizeof(int64_t) * world * cap_src);
  B.meta_slot_bytes = align_up(sizeof(int64_t) * world * 2);
  B.flags_slot_bytes = align_up(sizeof(unsigned long long) * 3 * world);
  B.off_ids_in = take(B.ids_slot_bytes * kPrepSlots);
  B.off_meta_in = take(B.meta_slot_bytes * kPrepSlots);
  B.off_meta_back = take(sizeof(int64_t) * world * 2);
  B.off_flags = take(B.flags_slot_bytes * kPrepSlots);
  B.off_uvals = take(sizeof(float) * cap_ids * d);
  B.off_grads_in = take(sizeof(float) * cap_recv * d);
  B.total = off;
  return B;
}

struct PrepSlot {
  const int64_t* ids = nullptr;
  int64_t n = 0;
  char* ws = nullptr;          // sort plan + unique index + inverse map of this batch
  int64_t* ukeys = nullptr;    // [cap_ids]
  int64_t* local = nullptr;    // [cap_ids] row inside the owner's shard of each unique id
  int64_t* bounds = nullptr;   // [world + 1] first unique index of each owner
  int64_t* n_unique = nullptr;
};

// device-side tables of pointers into every rank's block
struct PeerTables {
  int64_t** ids_in;                 // [world] rank g's ids_in region for THIS rank
  int64_t** meta_in;                // [world] rank g's meta_in entry for this rank
  unsigned long long** flag_ids;    // [world] rank g's ids flag for this rank
  float** uvals;                    // [world] rank r's uvals
  int64_t** meta_back;              // [world] rank r's meta_back entry for this rank (as owner)
  unsigned long long** flag_rows;   // [world] rank r's rows flag for this rank (as owner)
  float** grads_in;                 // [world] rank g's grads_in
  unsigned long long** flag_grads;  // [world] rank g's gradient flag for this rank
};

}  // namespace

struct b200emb_shard {
  int world = 1, rank = 0, device = 0, optimizer = 0;
  int64_t height = 0, hs = 0, plan_h = 0, d = 0;
  int64_t cap_ids = 0, cap_src = 0, cap_recv = 0;
  float *shard = nullptr, *m1 = nullptr, *m2 = nullptr;
  BlockLayout L{};
  char* block = nullptr;                 // this rank's peer-visible block
  std::vector<char*> peer;               // every rank's block as seen from this GPU
  std::vector<bool> opened;              // peer[r] came from cudaIpcOpenMemHandle
  bool connected = false;
  // private device memory
  char* priv = nullptr;
  PeerTables T[kPrepSlots] = {};         // per prepare slot (the id buffers and the flags are per slot)
  int64_t* recv_local = nullptr;         // [cap_recv] compacted row ids received as owner
  int64_t* seg_start = nullptr;          // [world + 1]
  unsigned long long* seq = nullptr;     // [kPrepSlots] use count of each slot, advanced by k_send_ids
  unsigned int* done_ctr = nullptr;      // [3] last-CTA detection of k_send_ids / owner gather / local MergeAdd
  int32_t* status = nullptr;
  char* owner_ws = nullptr;
  size_t owner_ws_bytes = 0, prep_ws_bytes = 0;
  PrepSlot slots[kPrepSlots];
  cudaStream_t side = nullptr;           // owner-side sort plan
  cudaEvent_t ev_fork = nullptr, ev_plan = nullptr;
};

namespace {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- requester: unique local rows -> their owners ---------------------------------------------------------
// dst region of owner g = ids_in[g] (already offset to this rank's region).  The last CTA to finish publishes,
// per owner, {count, offset of the first of these rows in this rank's unique list} and the ids flag, and
// advances the slot's use count, which every later kernel of the step reads as the value to wait for / publish.
__global__ void __launch_bounds__(256)
k_send_ids(const int64_t* __restrict__ local, const int64_t* __restrict__ bounds, int world, int64_t cap_src,
           PeerTables T, unsigned long long* epoch, unsigned int* done_ctr, int32_t* status) {
  const unsigned long long step = *reinterpret_cast<volatile unsigned long long*>(epoch) + 1;
  const int64_t total = bounds[world];
  for (int64_t u = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; u < total;
       u += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int g = 0;
    while (g + 1 < world && u >= bounds[g + 1]) ++g;
    const int64_t k = u - bounds[g];
    if (k < cap_src) T.ids_in[g][k] = local[u];
    else atomicOr(status, B200EMB_STATUS_SHARD_OVERFLOW);
  }
  __syncthreads();
  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    __threadfence();                 // cumulative gpu-scope fence per CTA, after the barrier that collects its stores
    s_last = atomicAdd(done_ctr, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  if (static_cast<int>(threadIdx.x) < world) {   // one system-scope release per destination covers ids and meta
    const int g = threadIdx.x;
    const int64_t cnt = min(bounds[g + 1] - bounds[g], cap_src);
    T.meta_in[g][0] = cnt;
    T.meta_in[g][1] = bounds[g];
    st_release_sys(T.flag_ids[g], step);
  }
  if (threadIdx.x == 0) {
    *done_ctr = 0;
    *epoch = step;
  }
}

// ---- owner: wait for every source's ids; gather their rows straight into the requesters' buffers -----------
// Compact index k runs over source 0's rows, source 1's rows, ...: row k is read from the shard and stored at
// (source r's uvals) + (its offset + k - start[r]) * D.  On the way the ids are written to recv[k] (the list the
// update's sort plan needs), recv[n_recv .. cap_recv) = pad_id (an id outside the shard: sorts last, skipped by
// every consumer), and seg_start[] = the prefix.  Every CTA waits for the flags and derives the (tiny) prefix
// itself; the rows flags (with each requester's start and accepted count) leave from this same kernel.
template <int LPR, int NV, bool VEC>
__global__ void __launch_bounds__(kBlock, 4)
k_owner_gather(const float* __restrict__ table, int64_t height, int d, const int64_t* __restrict__ ids_in,
               const int64_t* __restrict__ meta_in, const unsigned long long* __restrict__ flag_ids,
               const unsigned long long* seq, int world, int64_t cap_src, int64_t cap_recv, int64_t pad_id,
               float* const* __restrict__ peer_uvals, int64_t* __restrict__ recv, int64_t* __restrict__ seg_start,
               SignalSpec sig, int32_t* status) {
  constexpr int EPV = Epv<float, VEC>::value;
  constexpr int kGroups = kBlock / LPR;
  constexpr int kRows = 4;
  __shared__ int64_t s_start[kMaxWorld + 1];
  __shared__ int64_t s_off[kMaxWorld];
  __shared__ float* s_dst[kMaxWorld];
  wait_for_flags(flag_ids, seq, world, status);
  if (threadIdx.x == 0) {
    int64_t run = 0;
    for (int r = 0; r < world; ++r) {
      s_start[r] = run;
      int64_t c = meta_in[2 * r];
      if (run + c > cap_recv) {   // receive capacity exceeded: drop the excess and say so
        c = cap_recv - run;
        if (blockIdx.x == 0) atomicOr(status, B200EMB_STATUS_SHARD_OVERFLOW);
      }
      run += c;
    }
    s_start[world] = run;
  }
  if (static_cast<int>(threadIdx.x) < world) {
    s_off[threadIdx.x] = meta_in[2 * threadIdx.x + 1];
    s_dst[threadIdx.x] = peer_uvals[threadIdx.x];
  }
  __syncthreads();
  if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) <= world) seg_start[threadIdx.x] = s_start[threadIdx.x];
  const int64_t total = s_start[world];
  const int l = threadIdx.x % LPR;
  const int64_t group = static_cast<int64_t>(blockIdx.x) * kGroups + threadIdx.x / LPR;
  const int64_t num_groups = static_cast<int64_t>(gridDim.x) * kGroups;
  // id and destination of compact index k (software-pipelined: the next batch's ids are requested while this
  // batch's rows are in flight, so the id -> row dependency costs one round trip per batch, not two)
  const auto locate = [&](int64_t k, int64_t& id, float*& dst) {
    id = pad_id;
    dst = nullptr;
    if (k < total) {
      int src = 0;
      while (src + 1 < world && k >= s_start[src + 1]) ++src;
      const int64_t j = k - s_start[src];
      id = ids_in[static_cast<int64_t>(src) * cap_src + j];
      dst = s_dst[src] + (s_off[src] + j) * d;
    }
  };
  int64_t nid[kRows];
  float* ndst[kRows];
#pragma unroll
  for (int r = 0; r < kRows; ++r) locate(group + r * num_groups, nid[r], ndst[r]);
  for (int64_t base = group; base < cap_recv; base += num_groups * kRows) {
    int64_t id[kRows];
    float* dst[kRows];
    bool copy[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int64_t k = base + r * num_groups;
      id[r] = nid[r];
      dst[r] = ndst[r];
      copy[r] = dst[r] != nullptr && id[r] >= 0 && id[r] < height;
      if (l == 0 && k < cap_recv) recv[k] = id[r];
    }
#pragma unroll
    for (int r = 0; r < kRows; ++r) locate(base + num_groups * kRows + r * num_groups, nid[r], ndst[r]);
    for (int col0 = 0; col0 < d; col0 += LPR * NV * EPV) {
      float v[kRows][NV][EPV];
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk) {
          const int col = col0 + (l + kk * LPR) * EPV;
          if (copy[r] && col < d) {
            load_vec<float, VEC>(table + id[r] * d + col, v[r][kk]);
          } else {
#pragma unroll
            for (int e = 0; e < EPV; ++e) v[r][kk][e] = 0.f;
          }
        }
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk) {
          const int col = col0 + (l + kk * LPR) * EPV;
          if (dst[r] != nullptr && col < d) store_vec_stream<float, VEC>(dst[r] + col, v[r][kk]);
        }
    }
  }
  signal_peers(sig.flag_dst, sig.payload_dst, sig.starts, sig.epoch, sig.done_ctr, sig.world);
}

// ---- producer side of a flag: optional payload {a[r], b[r+1]-b[r]}, then the flag, one thread per destination ----
__global__ void k_signal(int world, int64_t* const* __restrict__ payload_dst, const int64_t* __restrict__ starts,
                         unsigned long long* const* __restrict__ flag_dst, const unsigned long long* epoch) {
  const int r = threadIdx.x;
  if (r >= world) return;
  if (payload_dst) {
    payload_dst[r][0] = starts[r];                   // first compact index of rank r's rows on this owner
    payload_dst[r][1] = starts[r + 1] - starts[r];   // how many of them the owner accepted
  }
  st_release_sys(flag_dst[r], *epoch);
}

int fail(cudaError_t) { return B200EMB_E_CUDA; }

template <typename P>
P* carve(char*& cur, size_t count) {
  P* p = reinterpret_cast<P*>(cur);
  cur += align_up(sizeof(P) * count);
  return p;
}

}  // namespace

extern "C" {

int b200emb_shard_create(int world, int rank, int64_t height, int64_t d, int optimizer, float* shard,
                         float* moment1, float* moment2, int64_t max_ids_per_step, int64_t max_recv_rows,
                         b200emb_shard_t** out) {
  if (!out || world < 1 || world > kMaxWorld || rank < 0 || rank >= world || height <= 0 || d <= 0 || !shard ||
      max_ids_per_step <= 0 || max_recv_rows <= 0 || (optimizer != B200EMB_SHARD_SGD && optimizer != B200EMB_SHARD_ADAM) ||
      (optimizer == B200EMB_SHARD_ADAM && (!moment1 || !moment2)))
    return B200EMB_E_INVALID;
  const int64_t hs = ceil_div(height, world);
  const int64_t plan_h = world > 1 ? world * hs : height;
  if (plan_h >= 0xFFFFFFFFLL || max_ids_per_step >= (1LL << 30) || max_recv_rows >= (1LL << 30)) return B200EMB_E_UNSUPPORTED;
  b200emb_shard* sh = new (std::nothrow) b200emb_shard();
  if (!sh) return B200EMB_E_INVALID;
  sh->world = world; sh->rank = rank; sh->optimizer = optimizer;
  sh->height = height; sh->hs = hs; sh->plan_h = plan_h; sh->d = d;
  sh->shard = shard; sh->m1 = moment1; sh->m2 = moment2;
  sh->cap_ids = max_ids_per_step;
  sh->cap_src = std::min(max_ids_per_step, max_recv_rows);   // one source can never fill more than the whole list
  sh->cap_recv = max_recv_rows;
  cudaGetDevice(&sh->device);
  sh->L = make_block(world, sh->cap_src, sh->cap_ids, sh->cap_recv, d);
  sh->peer.assign(world, nullptr);
  sh->opened.assign(world, false);
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&sh->block), sh->L.total);
  if (e == cudaSuccess) e = cudaMemset(sh->block, 0, sh->L.total);
  if (e == cudaSuccess) {
    const long long hdr[8] = {kBlockMagic, world, sh->cap_src, sh->cap_ids, sh->cap_recv, d, 0, 0};
    e = cudaMemcpy(sh->block + sh->L.off_header, hdr, sizeof(hdr), cudaMemcpyHostToDevice);
  }
  // private memory: tables, compact list, per-slot plan outputs, workspaces
  // a workspace good for ANY n <= n_max: the sort's tile size depends on n, so take the larger of the two regimes
  const auto ws_upto = [&](int64_t n_max) {
    return std::max(make_workspace(n_max, d, 4).total,
                    make_workspace(std::min<int64_t>(n_max, 148LL * 4 * 2 * 1024), d, 4).total);
  };
  sh->prep_ws_bytes = ws_upto(sh->cap_ids);
  sh->owner_ws_bytes = ws_upto(sh->cap_recv);
  const size_t w = static_cast<size_t>(world);
  size_t need = kPrepSlots * 8 * align_up(sizeof(void*) * w) + align_up(sizeof(int64_t) * sh->cap_recv) +
                align_up(sizeof(int64_t) * (w + 1)) + 3 * 256 +
                align_up(sh->owner_ws_bytes) +
                kPrepSlots * (align_up(sh->prep_ws_bytes) + 2 * align_up(sizeof(int64_t) * sh->cap_ids) +
                              align_up(sizeof(int64_t) * (w + 1)) + 256);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&sh->priv), need);
  if (e == cudaSuccess) e = cudaMemset(sh->priv, 0, need);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&sh->side, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sh->ev_fork, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&sh->ev_plan, cudaEventDisableTiming);
  if (e != cudaSuccess) {
    cudaGetLastError();
    b200emb_shard_destroy(sh);
    return B200EMB_E_CUDA;
  }
  char* cur = sh->priv;
  for (int t = 0; t < kPrepSlots; ++t) {
    sh->T[t].ids_in = carve<int64_t*>(cur, w);
    sh->T[t].meta_in = carve<int64_t*>(cur, w);
    sh->T[t].flag_ids = carve<unsigned long long*>(cur, w);
    sh->T[t].uvals = carve<float*>(cur, w);
    sh->T[t].meta_back = carve<int64_t*>(cur, w);
    sh->T[t].flag_rows = carve<unsigned long long*>(cur, w);
    sh->T[t].grads_in = carve<float*>(cur, w);
    sh->T[t].flag_grads = carve<unsigned long long*>(cur, w);
  }
  sh->recv_local = carve<int64_t>(cur, sh->cap_recv);
  sh->seg_start = carve<int64_t>(cur, w + 1);
  sh->seq = carve<unsigned long long>(cur, kPrepSlots);
  sh->done_ctr = carve<unsigned int>(cur, 3);
  sh->status = carve<int32_t>(cur, 1);
  sh->owner_ws = cur; cur += align_up(sh->owner_ws_bytes);
  for (int s = 0; s < kPrepSlots; ++s) {
    PrepSlot& p = sh->slots[s];
    p.ws = cur; cur += align_up(sh->prep_ws_bytes);
    p.ukeys = carve<int64_t>(cur, sh->cap_ids);
    p.local = carve<int64_t>(cur, sh->cap_ids);
    p.bounds = carve<int64_t>(cur, w + 1);
    p.n_unique = carve<int64_t>(cur, 1);
  }
  *out = sh;
  return B200EMB_OK;
}

int b200emb_shard_export(b200emb_shard_t* sh, unsigned char* handle64) {
  if (!sh || !handle64 || !sh->block) return B200EMB_E_INVALID;
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, sh->block) != cudaSuccess) { cudaGetLastError(); return B200EMB_E_CUDA; }
  std::memcpy(handle64, &h, 64);
  return B200EMB_OK;
}

void* b200emb_shard_block(b200emb_shard_t* sh) { return sh ? sh->block : nullptr; }

static int finish_connect(b200emb_shard* sh) {
  const int w = sh->world, me = sh->rank;
  const BlockLayout& L = sh->L;
  for (int g = 0; g < w; ++g) {   // pointers into a peer's block are computed with THIS rank's layout: it must match
    long long hdr[8] = {0};
    if (cudaMemcpy(hdr, sh->peer[g] + L.off_header, sizeof(hdr), cudaMemcpyDeviceToHost) != cudaSuccess) {
      cudaGetLastError();
      return B200EMB_E_CUDA;
    }
    if (hdr[0] != kBlockMagic || hdr[1] != w || hdr[2] != sh->cap_src || hdr[3] != sh->cap_ids ||
        hdr[4] != sh->cap_recv || hdr[5] != sh->d)
      return B200EMB_E_INVALID;   // create() was called with different sizes on that rank
  }
  std::vector<void*> h(8 * static_cast<size_t>(w));
  for (int t = 0; t < kPrepSlots; ++t) {
    for (int g = 0; g < w; ++g) {
      char* pb = sh->peer[g];
      char* fl = pb + L.off_flags + L.flags_slot_bytes * t;
      h[0 * w + g] = pb + L.off_ids_in + L.ids_slot_bytes * t + sizeof(int64_t) * static_cast<size_t>(me) * sh->cap_src;
      h[1 * w + g] = pb + L.off_meta_in + L.meta_slot_bytes * t + sizeof(int64_t) * 2 * me;
      h[2 * w + g] = fl + sizeof(unsigned long long) * (0 * w + me);
      h[3 * w + g] = pb + L.off_uvals;
      h[4 * w + g] = pb + L.off_meta_back + sizeof(int64_t) * 2 * me;
      h[5 * w + g] = fl + sizeof(unsigned long long) * (1 * w + me);
      h[6 * w + g] = pb + L.off_grads_in;
      h[7 * w + g] = fl + sizeof(unsigned long long) * (2 * w + me);
    }
    void* dst[8] = {sh->T[t].ids_in, sh->T[t].meta_in, sh->T[t].flag_ids, sh->T[t].uvals,
                    sh->T[t].meta_back, sh->T[t].flag_rows, sh->T[t].grads_in, sh->T[t].flag_grads};
    for (int k = 0; k < 8; ++k)
      if (cudaMemcpy(dst[k], &h[static_cast<size_t>(k) * w], sizeof(void*) * w, cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaGetLastError();
        return B200EMB_E_CUDA;
      }
  }
  sh->connected = true;
  return B200EMB_OK;
}

int b200emb_shard_connect(b200emb_shard_t* sh, const unsigned char* handles) {
  if (!sh || !handles) return B200EMB_E_INVALID;
  for (int r = 0; r < sh->world; ++r) {
    if (r == sh->rank) { sh->peer[r] = sh->block; continue; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handles + 64 * static_cast<size_t>(r), 64);
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return B200EMB_E_CUDA; }
    sh->peer[r] = static_cast<char*>(p);
    sh->opened[r] = true;
  }
  return finish_connect(sh);
}

int b200emb_shard_connect_ptrs(b200emb_shard_t* sh, void* const* blocks) {
  if (!sh || !blocks) return B200EMB_E_INVALID;
  for (int r = 0; r < sh->world; ++r) {
    if (!blocks[r]) return B200EMB_E_INVALID;
    sh->peer[r] = static_cast<char*>(blocks[r]);
  }
  if (sh->peer[sh->rank] != sh->block) return B200EMB_E_INVALID;
  return finish_connect(sh);
}

int b200emb_shard_prepare(b200emb_shard_t* sh, int slot, const int64_t* ids, int64_t n, void* stream) {
  if (!sh || slot < 0 || slot >= kPrepSlots || n < 0 || n > sh->cap_ids || (n > 0 && !ids)) return B200EMB_E_INVALID;
  if (!sh->connected) return B200EMB_E_INVALID;
  PrepSlot& p = sh->slots[slot];
  p.ids = ids;
  p.n = n;
  const int rc = b200emb_plan_unique_rows(ids, n, sh->height, sh->world, p.ukeys, p.n_unique, p.bounds, nullptr,
                                          p.local, p.ws, sh->prep_ws_bytes, sh->status, stream);
  if (rc != B200EMB_OK) return rc;
  // the id exchange needs the ids only, too: the owners find them waiting when the step starts
  const int grid = static_cast<int>(std::min<int64_t>(std::max<int64_t>(ceil_div(n, 256), 1), 148 * 4));
  k_send_ids<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p.local, p.bounds, sh->world, sh->cap_src, sh->T[slot],
                                                                 sh->seq + slot, sh->done_ctr, sh->status);
  return cudaGetLastError() == cudaSuccess ? B200EMB_OK : B200EMB_E_CUDA;
}

// phases of lookup: 0 (nothing: the ids went out in prepare) | 1 owner gather + flags | 2 fork owner sort plan | 3 expand
int b200emb_shard_lookup_phases(b200emb_shard_t* sh, int slot, float* out, void* stream, int first, int last) {
  if (!sh || !sh->connected || slot < 0 || slot >= kPrepSlots || first < 0 || last > 3 || first > last) return B200EMB_E_INVALID;
  PrepSlot& p = sh->slots[slot];
  if (p.n > 0 && !out) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int w = sh->world;
  const BlockLayout& L = sh->L;
  unsigned long long* flags = reinterpret_cast<unsigned long long*>(sh->block + L.off_flags + L.flags_slot_bytes * slot);
  const unsigned long long* seq = sh->seq + slot;
  cudaError_t e = cudaSuccess;
  if (first <= 1 && 1 <= last) {
    // the rows and the rows flags leave from the same kernel: every CTA fences its peer stores before the last
    // one raises the flags (and tells each requester where its rows sit in this owner's compact list)
    SignalSpec sg;
    sg.flag_dst = sh->T[slot].flag_rows;
    sg.payload_dst = sh->T[slot].meta_back;
    sg.starts = sh->seg_start;
    sg.epoch = seq;
    sg.done_ctr = sh->done_ctr + 1;
    sg.world = w;
    const int64_t* ids_in = reinterpret_cast<const int64_t*>(sh->block + L.off_ids_in + L.ids_slot_bytes * slot);
    const int64_t* meta_in = reinterpret_cast<const int64_t*>(sh->block + L.off_meta_in + L.meta_slot_bytes * slot);
    const bool aligned = is_aligned16(sh->shard) && (sh->d % 4) == 0;
    dispatch_row_shape<float>(sh->d, aligned, [&]<int LPR, int NV, bool VEC>() {
      constexpr int kGroups = kBlock / LPR;
      const int grid = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(ceil_div(sh->cap_recv, kGroups), 148 * 4)));
      k_owner_gather<LPR, NV, VEC><<<grid, kBlock, 0, s>>>(sh->shard, sh->hs, static_cast<int>(sh->d), ids_in, meta_in,
                                                          flags, seq, w, sh->cap_src, sh->cap_recv, sh->hs,
                                                          sh->T[slot].uvals, sh->recv_local, sh->seg_start, sg, sh->status);
    });
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(e);
  }
  if (first <= 2 && 2 <= last) {
    // the update's sort plan needs the received ids only: build it beside the expansion and the local MergeAdd
    if ((e = cudaEventRecord(sh->ev_fork, s)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamWaitEvent(sh->side, sh->ev_fork, 0)) != cudaSuccess) return fail(e);
    const Workspace OW = make_workspace(sh->cap_recv, sh->d, 4);
    e = launch_plan(sh->recv_local, sh->cap_recv, sh->hs, sh->owner_ws, OW, nullptr, sh->side, nullptr);
    if (e != cudaSuccess) return fail(e);
    if ((e = cudaEventRecord(sh->ev_plan, sh->side)) != cudaSuccess) return fail(e);
  }
  if (first <= 3 && 3 <= last && p.n > 0) {
    const Workspace PW = make_workspace(p.n, 1, 4);
    SortedPlan plan = plan_view(p.n, sh->plan_h, p.ws, PW);
    plan.perm = reinterpret_cast<const uint32_t*>(p.ws + PW.off_perm[(plan.passes & 1) ^ 1]);   // the inverse map
    WaitSpec wt;
    wt.flags = flags + w;
    wt.epoch = seq;
    wt.world = w;
    wt.status = sh->status;
    e = launch_permute_rows<float>(reinterpret_cast<const float*>(sh->block + L.off_uvals), plan,
                                   reinterpret_cast<const uint32_t*>(p.ws + PW.off_uidx), p.n, sh->d, out, 0, s, wt);
    if (e != cudaSuccess) return fail(e);
  }
  return B200EMB_OK;
}

int b200emb_shard_lookup(b200emb_shard_t* sh, int slot, float* out, void* stream) {
  return b200emb_shard_lookup_phases(sh, slot, out, stream, 0, 3);
}

// phases of update: 0 local MergeAdd into the owners' buffers + signal | 1 owner fused update
int b200emb_shard_update_phases(b200emb_shard_t* sh, int slot, const float* d_out, const float* lr,
                                const float* beta1_pow, const float* beta2_pow, float beta1, float beta2,
                                float epsilon, void* stream, int first, int last) {
  if (!sh || !sh->connected || slot < 0 || slot >= kPrepSlots || !lr || first < 0 || last > 1 || first > last)
    return B200EMB_E_INVALID;
  if (sh->optimizer == B200EMB_SHARD_ADAM && (!beta1_pow || !beta2_pow)) return B200EMB_E_INVALID;
  PrepSlot& p = sh->slots[slot];
  if (p.n > 0 && !d_out) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int w = sh->world;
  const BlockLayout& L = sh->L;
  unsigned long long* flags = reinterpret_cast<unsigned long long*>(sh->block + L.off_flags + L.flags_slot_bytes * slot);
  const unsigned long long* seq = sh->seq + slot;
  cudaError_t e = cudaSuccess;
  if (first <= 0 && 0 <= last) {
    if (p.n > 0) {
      const Workspace PW = make_workspace(p.n, sh->d, 4);
      UpdateArgs<float> a;
      a.op = UpdateOp::kMergeAdd;
      a.param = const_cast<float*>(d_out);   // unused: every merged row goes through the scatter map
      a.uidx = reinterpret_cast<const uint32_t*>(p.ws + PW.off_uidx);
      a.shard_bounds = p.bounds;
      a.shard_dst = sh->T[slot].grads_in;
      a.shard_meta = reinterpret_cast<const int64_t*>(sh->block + L.off_meta_back);
      a.shards = w;
      a.dst_aligned16 = (sh->d % 4) == 0;
      a.signal.flag_dst = sh->T[slot].flag_grads;      // merged rows and their flag leave from the same kernel
      a.signal.epoch = seq;
      a.signal.done_ctr = sh->done_ctr + 2;
      a.signal.world = w;
      const SortedPlan plan = plan_view(p.n, sh->plan_h, p.ws, PW);
      e = launch_update<float>(plan, d_out, p.n, sh->d, sh->plan_h, a, false, p.ws, PW, sh->status, s);
      if (e != cudaSuccess) return fail(e);
    } else {   // an empty batch still has to tell the owners that nothing is coming
      k_signal<<<1, ((w + 31) / 32) * 32, 0, s>>>(w, nullptr, nullptr, sh->T[slot].flag_grads, seq);
      if ((e = cudaGetLastError()) != cudaSuccess) return fail(e);
    }
  }
  if (first <= 1 && 1 <= last) {
    if ((e = cudaStreamWaitEvent(s, sh->ev_plan, 0)) != cudaSuccess) return fail(e);
    const Workspace OW = make_workspace(sh->cap_recv, sh->d, 4);
    UpdateArgs<float> a;
    a.op = sh->optimizer == B200EMB_SHARD_ADAM ? UpdateOp::kAdam : UpdateOp::kSgd;
    a.param = sh->shard;
    a.moment1 = sh->m1;
    a.moment2 = sh->m2;
    a.lr = lr;
    a.beta1_pow = beta1_pow;
    a.beta2_pow = beta2_pow;
    a.beta1 = beta1; a.beta2 = beta2; a.epsilon = epsilon;
    a.wait.flags = flags + 2 * w;
    a.wait.epoch = seq;
    a.wait.world = w;
    a.wait.status = sh->status;
    const SortedPlan plan = plan_view(sh->cap_recv, sh->hs, sh->owner_ws, OW);
    e = launch_update<float>(plan, reinterpret_cast<const float*>(sh->block + L.off_grads_in), sh->cap_recv, sh->d,
                             sh->hs, a, false, sh->owner_ws, OW, sh->status, s);
    if (e != cudaSuccess) return fail(e);
  }
  return B200EMB_OK;
}

int b200emb_shard_update(b200emb_shard_t* sh, int slot, const float* d_out, const float* lr, const float* beta1_pow,
                         const float* beta2_pow, float beta1, float beta2, float epsilon, void* stream) {
  return b200emb_shard_update_phases(sh, slot, d_out, lr, beta1_pow, beta2_pow, beta1, beta2, epsilon, stream, 0, 1);
}

int32_t* b200emb_shard_status_ptr(b200emb_shard_t* sh) { return sh ? sh->status : nullptr; }

int b200emb_shard_info(b200emb_shard_t* sh, int64_t* out8) {
  if (!sh || !out8) return B200EMB_E_INVALID;
  out8[0] = sh->world; out8[1] = sh->rank; out8[2] = sh->hs; out8[3] = sh->plan_h;
  out8[4] = sh->cap_ids; out8[5] = sh->cap_src; out8[6] = sh->cap_recv; out8[7] = static_cast<int64_t>(sh->L.total);
  return B200EMB_OK;
}

// device pointers a caller may want to look at after a synchronize (debugging, tests, byte accounting)
int b200emb_shard_debug_ptrs(b200emb_shard_t* sh, int slot, void** out6) {
  if (!sh || !out6 || slot < 0 || slot >= kPrepSlots) return B200EMB_E_INVALID;
  out6[0] = sh->slots[slot].bounds;
  out6[1] = sh->slots[slot].n_unique;
  out6[2] = sh->seg_start;
  out6[3] = sh->recv_local;
  out6[4] = sh->seq + slot;
  out6[5] = sh->status;
  return B200EMB_OK;
}

int b200emb_shard_destroy(b200emb_shard_t* sh) {
  if (!sh) return B200EMB_OK;
  for (int r = 0; r < sh->world; ++r)
    if (r < static_cast<int>(sh->opened.size()) && sh->opened[r] && sh->peer[r]) cudaIpcCloseMemHandle(sh->peer[r]);
  if (sh->side) cudaStreamDestroy(sh->side);
  if (sh->ev_fork) cudaEventDestroy(sh->ev_fork);
  if (sh->ev_plan) cudaEventDestroy(sh->ev_plan);
  if (sh->priv) cudaFree(sh->priv);
  if (sh->block) cudaFree(sh->block);
  cudaGetLastError();
  delete sh;
  return B200EMB_OK;
}

}  // extern "C"
