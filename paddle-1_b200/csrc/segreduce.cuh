// segreduce.cuh — duplicate-row reduction fused with the consumer (K3/K4/K5a/K5b), sm_100a.
//
// Input: the sort plan (sorted ids + original positions) and the [N,D] gradient rows.
// One skeleton, four consumers selected at compile time:
//   kMergeAdd     out_value[u,:]  = sum           scatter::MergeAdd (selected_rows_functor.cc:193-228)
//   kAddToTensor  tensor[id,:]   += sum           SelectedRowsAddToTensor (:153-178)
//   kSgd          param[id,:]    -= lr*sum        sparse SGD (sgd_op.h:47-81)
//   kAdam         lazy Adam on (param,m1,m2)[id]  SparseAdamFunctor (adam_op.h:199-219)
//   kAdagrad      moment += sum^2; param -= lr*sum/(sqrt(moment)+eps)   SparseAdagradFunctor (adagrad_op.cc:108-141)
// so merged gradient rows never travel through HBM: each input row is read once, each touched
// parameter row is read and written once.
//
// FAST mode (k_chunk_reduce): the sorted element range is cut into equal chunks, one
// per lane group, so work is balanced no matter how skewed the ids are (a Zipf head id owns
// 10-25% of all rows).  A group walks its chunk through a shared-memory ring filled by cp.async
// (LDGSTS): every lane copies ITS 16 bytes of each gradient row — and of the parameter / moment rows of
// every segment that closes at that row — into slots only it reads back, so the ring needs no barrier at
// all, only cp.async.wait_group; three batches are in flight while one is consumed (the loads of batch
// k+3 are issued as soon as batch k has been added up), and no payload lives in registers.
// Segments that close inside the chunk are consumed immediately.  Segments cut by a chunk
// boundary leave partial sums in shared memory, are combined across the CTA's groups in fixed
// order, and — if they also cross the CTA boundary — leave one carry row per CTA; the last CTA
// of such a chain to finish (an integer arrival counter, nobody waits) sums the chain's carries in
// a fixed order.  No floating-point atomics anywhere: the summation tree is fixed by (N, D) alone,
// hence run-to-run deterministic.
//
// DETERMINISTIC mode (k_segment_det): one lane group per unique row walks its duplicates in input
// order (the stable sort keeps it) with separately rounded IEEE operations, reproducing the
// reference CPU loops bit for bit.  Its critical path is the longest duplicate run; it is the
// parity mode, not the speed mode.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "internal.h"

namespace b200emb {
namespace {

constexpr uint32_t kHead = 1u, kTail = 2u, kThrough = 4u;             // per-group flags

// ---- developer trace (tools/ only: built with -DB200EMB_TRACE) -------------------------------------
#ifdef B200EMB_TRACE
__device__ unsigned long long g_trace[8192][8];
#define B200EMB_STAMP(slot)                                                                   \
  do {                                                                                        \
    if (threadIdx.x == 0 && blockIdx.x < 8192) {                                              \
      unsigned long long t_;                                                                  \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                  \
      g_trace[blockIdx.x][slot] = t_;                                                         \
    }                                                                                         \
  } while (0)
#else
#define B200EMB_STAMP(slot) do {} while (0)
#endif

template <typename T>
struct Params {
  T* param; T* m1; T* m2;
  const T* lr; const T* b1p; const T* b2p;
  T beta1, beta2, eps;
  const uint32_t* uidx;
  int d;
  uint32_t height;
  // kMergeAdd with a scatter map: merged row u goes to shard_dst[g] + (u - shard_bounds[g]) * d for the
  // shard g with shard_bounds[g] <= u < shard_bounds[g+1].  The destinations may be peer-GPU memory
  // (NVLink stores), which fuses the local pre-reduction with the gradient all-to-all.
  const int64_t* shard_bounds;
  T* const* shard_dst;
  int shards;
  const int64_t* shard_meta;  // optional pairs {rows to skip at the start of destination g, rows it accepts}
  WaitSpec wait;              // sharded owner side: `value` is stored by other GPUs; wait for their flags first
  SignalSpec signal;          // sharded requester side: merged rows go to other GPUs; publish when all are stored
};

template <typename T>
struct Carry {
  uint32_t* arrive;   // [ctas] arrivals at the chain that STARTS in this CTA; zero between launches
  T* head; T* tail;   // [ctas][d] partial sums of the segment entering / leaving each CTA
};

// Largest b in [lo, hi) with pred(b) true, for a predicate that is true on [lo, b] and false on
// (b, hi] (pred(lo) true and pred(hi) false are known, not evaluated).  One warp, 32 probes per
// memory round trip: a chain of any length is bracketed in <= 4 round trips.
template <class Pred>
__device__ __forceinline__ int warp_last_true(int lo, int hi, int lane, Pred pred) {
  while (hi - lo > 1) {
    const int step = (hi - lo + 31) / 32;
    const int p = lo + step * (lane + 1);
    const bool t = p < hi ? pred(p) : false;
    const uint32_t f = __ballot_sync(0xffffffffu, !t);   // lane 31 probes >= hi: never empty
    const int i = __ffs(f) - 1;
    hi = min(hi, lo + step * (i + 1));
    lo = lo + step * i;
  }
  return lo;
}

template <typename T, int NV, int EPV>
__device__ __forceinline__ void zero_acc(T (&a)[NV][EPV]) {
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int e = 0; e < EPV; ++e) a[k][e] = T(0);
}

// The consumer of one fully reduced row is split in two so that callers can issue the state
// loads (parameter / moment rows) of several rows together with the gradient-row loads and pay
// one memory round trip per batch instead of one per row.
template <UpdateOp OP> struct NumState {
  static constexpr int value =
      OP == UpdateOp::kAdam ? 3 : (OP == UpdateOp::kAdagrad ? 2 : (OP == UpdateOp::kMergeAdd ? 0 : 1));
};
template <typename T, UpdateOp OP, int NV, int EPV>
struct RowState { T v[NumState<OP>::value == 0 ? 1 : NumState<OP>::value][NV][EPV]; };

// ---- cp.async (LDGSTS) ring helpers ----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// BYTES = 16: L2 only (.cg, streaming data); 4 / 8 (scalar path of odd widths): .ca is the only form
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc) {
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_addr(smem_dst)), "l"(gsrc) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" :: "r"(smem_addr(smem_dst)), "l"(gsrc), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

constexpr int kBoundarySlack = 64;        // a CTA boundary moves forward by at most this many elements

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void load_state(const Params<T>& P, uint32_t key, int col0, int l,
                                           RowState<T, OP, NV, Epv<T, VEC>::value>& st) {
  constexpr int EPV = Epv<T, VEC>::value;
  if constexpr (OP == UpdateOp::kMergeAdd) return;
  if (key >= P.height) return;
  const int64_t base = static_cast<int64_t>(key) * P.d;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int col = col0 + (l + k * LPR) * EPV;
    if (col >= P.d) continue;
    load_vec<T, VEC>(P.param + base + col, st.v[0][k]);
    if constexpr (OP == UpdateOp::kAdam) {
      load_vec<T, VEC>(P.m1 + base + col, st.v[1][k]);
      load_vec<T, VEC>(P.m2 + base + col, st.v[2][k]);
    }
    if constexpr (OP == UpdateOp::kAdagrad) load_vec<T, VEC>(P.m1 + base + col, st.v[1][k]);
  }
}

// `eref` = any sorted position inside the segment (kMergeAdd looks its output row up there).
template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void apply_state(const Params<T>& P, uint32_t key, uint32_t eref, int col0,
                                            int l, const T (&acc)[NV][Epv<T, VEC>::value],
                                            RowState<T, OP, NV, Epv<T, VEC>::value>& st,
                                            T* merge_dst_known = nullptr, bool merge_dst_is_known = false) {
  constexpr int EPV = Epv<T, VEC>::value;
  using A = Arith<T>;
  if (key >= P.height) return;  // out-of-range row: flagged by the sort plan, skipped here
  if (OP == UpdateOp::kMergeAdd && merge_dst_is_known && merge_dst_known == nullptr) return;   // refused by its owner
  const int64_t row = (OP == UpdateOp::kMergeAdd) ? (merge_dst_is_known ? 0 : static_cast<int64_t>(P.uidx[eref]))
                                                  : static_cast<int64_t>(key);
  const int64_t base = row * P.d;
  T* merge_dst = merge_dst_is_known ? merge_dst_known : P.param + base;
  if constexpr (OP == UpdateOp::kMergeAdd) {
    if (P.shards > 0 && !merge_dst_is_known) {
      int g = 0;
      while (g + 1 < P.shards && row >= P.shard_bounds[g + 1]) ++g;
      int64_t k = row - P.shard_bounds[g];
      if (P.shard_meta) {
        if (k >= P.shard_meta[2 * g + 1]) return;   // the owner ran out of room (it raised the overflow status bit)
        k += P.shard_meta[2 * g];
      }
      merge_dst = P.shard_dst[g] + k * P.d;
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int col = col0 + (l + k * LPR) * EPV;
    if (col >= P.d) continue;
    if constexpr (OP == UpdateOp::kMergeAdd) {
      store_vec<T, VEC>(merge_dst + col, acc[k]);
    } else if constexpr (OP == UpdateOp::kAddToTensor) {
#pragma unroll
      for (int e = 0; e < EPV; ++e) st.v[0][k][e] = A::add(st.v[0][k][e], acc[k][e]);
      store_vec<T, VEC>(P.param + base + col, st.v[0][k]);
    } else if constexpr (OP == UpdateOp::kSgd) {
      const T lr = *P.lr;
#pragma unroll
      for (int e = 0; e < EPV; ++e) st.v[0][k][e] = A::sub(st.v[0][k][e], A::mul(lr, acc[k][e]));
      store_vec<T, VEC>(P.param + base + col, st.v[0][k]);
    } else if constexpr (OP == UpdateOp::kAdagrad) {
      // adagrad_op.cc:119-138: moment += g*g (scatter::Mul + SelectedRowsAddToTensor), then
      // param -= lr*g / (sqrt(moment) + eps), evaluated left to right
      const T lr = *P.lr;
#pragma unroll
      for (int e = 0; e < EPV; ++e) {
        const T g = acc[k][e];
        const T m = A::add(st.v[1][k][e], A::mul(g, g));
        st.v[1][k][e] = m;
        st.v[0][k][e] = A::sub(st.v[0][k][e], A::div(A::mul(lr, g), A::add(A::sqrt(m), P.eps)));
      }
      store_vec<T, VEC>(P.m1 + base + col, st.v[1][k]);
      store_vec<T, VEC>(P.param + base + col, st.v[0][k]);
    } else {  // kAdam — operation order of adam_op.h:209-213, every step rounded in T
      const T one = T(1);
      const T lr_t = A::mul(*P.lr, A::div(A::sqrt(A::sub(one, *P.b2p)), A::sub(one, *P.b1p)));
      const T omb1 = A::sub(one, P.beta1), omb2 = A::sub(one, P.beta2);
#pragma unroll
      for (int e = 0; e < EPV; ++e) {
        const T g = acc[k][e];
        const T m1 = A::add(A::mul(P.beta1, st.v[1][k][e]), A::mul(omb1, g));
        const T m2 = A::add(A::mul(P.beta2, st.v[2][k][e]), A::mul(A::mul(omb2, g), g));
        st.v[0][k][e] = A::sub(st.v[0][k][e], A::mul(lr_t, A::div(m1, A::add(A::sqrt(m2), P.eps))));
        st.v[1][k][e] = m1;
        st.v[2][k][e] = m2;
      }
      store_vec<T, VEC>(P.m1 + base + col, st.v[1][k]);
      store_vec<T, VEC>(P.m2 + base + col, st.v[2][k]);
      store_vec<T, VEC>(P.param + base + col, st.v[0][k]);
    }
  }
}

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void consume(const Params<T>& P, uint32_t key, uint32_t eref, int col0,
                                        int l, const T (&acc)[NV][Epv<T, VEC>::value]) {
  RowState<T, OP, NV, Epv<T, VEC>::value> st;
  load_state<T, OP, LPR, NV, VEC>(P, key, col0, l, st);
  apply_state<T, OP, LPR, NV, VEC>(P, key, eref, col0, l, acc, st);
}

// Closes one chain: the segment that leaves CTA `c` (tail carry) and enters the next `len` CTAs
// (head carries).  Run by the LAST CTA of the chain to arrive, with the whole CTA: the lane
// groups sum the head rows strided by group with all loads in flight, group 0 adds the per-group
// partials to the tail in group order — the summation order depends on (N, D) alone, not on
// which CTA happens to arrive last.
template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void close_chain(int c, int len, uint32_t key, uint32_t eref,
                                            const Params<T>& P, const Carry<T>& C, T* s_part) {
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int GPC = kBlock / LPR;
  constexpr int COLS = LPR * NV * EPV;
  constexpr int kRows = NV == 1 ? 8 : 4;  // head rows in flight per group
  using A = Arith<T>;
  const int tid = threadIdx.x, g = tid / LPR, l = tid % LPR;
  const int d = P.d;
  for (int col0 = 0; col0 < d; col0 += COLS) {
    T sum[NV][EPV];
    RowState<T, OP, NV, EPV> st;
    if (g == 0) {  // tail row and the row's state: issued before the head rows so all these loads overlap
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int col = col0 + (l + k * LPR) * EPV;
        if (col < d) load_vec_cg<T, VEC>(C.tail + static_cast<int64_t>(c) * d + col, sum[k]);
        else {
#pragma unroll
          for (int e = 0; e < EPV; ++e) sum[k][e] = T(0);
        }
      }
      load_state<T, OP, LPR, NV, VEC>(P, key, col0, l, st);
    }
    T part[NV][EPV];
    zero_acc<T, NV, EPV>(part);
    for (int r0 = g; r0 < len; r0 += GPC * kRows) {
      T h[kRows][NV][EPV];
#pragma unroll
      for (int q = 0; q < kRows; ++q) {
        const int r = r0 + q * GPC;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          if (r < len && col < d)
            load_vec_cg<T, VEC>(C.head + static_cast<int64_t>(c + 1 + r) * d + col, h[q][k]);
          else {
#pragma unroll
            for (int e = 0; e < EPV; ++e) h[q][k][e] = T(0);
          }
        }
      }
#pragma unroll
      for (int q = 0; q < kRows; ++q)
        if (r0 + q * GPC < len) {
#pragma unroll
          for (int k = 0; k < NV; ++k)
#pragma unroll
            for (int e = 0; e < EPV; ++e) part[k][e] = A::add(part[k][e], h[q][k][e]);
        }
    }
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int e = 0; e < EPV; ++e) s_part[g * COLS + (l + k * LPR) * EPV + e] = part[k][e];
    __syncthreads();
    if (g == 0) {
      const int used = min(len, GPC);
      for (int j = 0; j < used; ++j)
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int e = 0; e < EPV; ++e)
            sum[k][e] = A::add(sum[k][e], s_part[j * COLS + (l + k * LPR) * EPV + e]);
      apply_state<T, OP, LPR, NV, VEC>(P, key, eref, col0, l, sum, st);
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Geometry of one launch, fixed by (N, D) alone (never by the consumer): `chunk` sorted elements per lane
// group nominally, CTA c nominally owns [c*stride, (c+1)*stride) with stride = groups*chunk.  A nominal
// boundary that falls inside a segment is moved FORWARD to that segment's end when the end is at most
// `slack` elements away, so only segments longer than that (the Zipf head) are ever shared between CTAs;
// a CTA then splits its adjusted range evenly over its groups.
struct ChunkGeom {
  int n, chunk, slack;
};

// The shared-memory ring holds kRingStages batches of kBatch rows; kBatch is a compile-time property of the
// row shape (a launch-time batch with predicated, fully unrolled loops cost the one-row-per-batch Adam shapes
// 14 %), sized so that the ring of a CTA stays near 48 KB and three CTAs fit an SM for the common shapes.
constexpr int kRingStages = 3;
constexpr int kRingTargetBytes = 48 * 1024;
template <typename T, UpdateOp OP, int NV, bool VEC>
struct Ring {
  static constexpr int kVB = Epv<T, VEC>::value * static_cast<int>(sizeof(T));   // bytes per lane per vector
  // state rows ride in the ring beside the gradient row while a slot stays <= 128 bytes per thread; wider
  // shapes (>= 512 floats with moments) read them with plain loads when the segment closes
  static constexpr bool kStateInRing = NumState<OP>::value > 0 && (1 + NumState<OP>::value) * NV <= 8;
  static constexpr int kParts = kStateInRing ? 1 + NumState<OP>::value : 1;      // gradient row + state rows
  static constexpr int kSlotBytes = kParts * NV * kVB * kBlock;                  // one slot, whole CTA
  // one more slot holds the state of the segment a group leaves open (single-state consumers only: for
  // Adam / Adagrad that slot would cost a resident CTA)
  static constexpr bool kTailInRing = kStateInRing && NumState<OP>::value == 1;
  static constexpr int kFit = kRingTargetBytes / (kRingStages * kSlotBytes);
  static constexpr int kBatch = kFit >= 4 ? 4 : (kFit >= 2 ? 2 : 1);
  static constexpr int kBytes = (kRingStages * kBatch + (kTailInRing ? 1 : 0)) * kSlotBytes;
};

__device__ __forceinline__ void lds_vec(uint32_t addr, float (&v)[4]) {
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(addr));
}
__device__ __forceinline__ void lds_vec(uint32_t addr, double (&v)[2]) {
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v[0]), "=d"(v[1]) : "r"(addr));
}
__device__ __forceinline__ void lds_vec(uint32_t addr, float (&v)[1]) {
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v[0]) : "r"(addr));
}
__device__ __forceinline__ void lds_vec(uint32_t addr, double (&v)[1]) {
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v[0]) : "r"(addr));
}
template <int BYTES>
__device__ __forceinline__ void cp_async_s(uint32_t smem_dst, const void* gsrc) {
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_dst), "l"(gsrc) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" :: "r"(smem_dst), "l"(gsrc), "n"(BYTES) : "memory");
}

constexpr uint32_t kPosMask = 0x3FFFFFFFu;   // N < 2^30: the two top bits of a staged position carry flags
constexpr uint32_t kCloses = 0x80000000u;    // the segment ends at this element
constexpr uint32_t kDirect = 0x40000000u;    // ... and started inside this group's chunk: consumed on the spot

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void chunk_reduce_body(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ perm,
                                                  const ChunkGeom& G, const T* __restrict__ value,
                                                  const Params<T>& P, const Carry<T>& C) {
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int GPC = kBlock / LPR;
  constexpr int COLS = LPR * NV * EPV;
  constexpr int NS = NumState<OP>::value;
  using RG = Ring<T, OP, NV, VEC>;
  constexpr int VB = RG::kVB, S = kRingStages;
  constexpr int B = RG::kBatch;
  constexpr uint32_t kVecStride = kBlock * VB;                 // between the vectors of one slot
  constexpr uint32_t kPartStride = NV * kVecStride;            // between the parts (row, state rows) of one slot
  constexpr uint32_t kSlotStride = RG::kSlotBytes;
  constexpr int kTailSlot = S * B;                                 // state of the segment a group leaves open
  using A = Arith<T>;
  const int n = G.n;
  const int stride = GPC * G.chunk;                            // nominal sorted elements per CTA
  const int win = stride + G.slack + 2;                        // staged key window
  const int tid = threadIdx.x, g = tid / LPR, l = tid % LPR;
  const int lane = tid & 31, warp = tid >> 5;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // ring[slot][part][vector][thread], filled by cp.async: a thread only ever touches its own column, so the
  // ring needs no barrier, only cp.async.wait_group
  const uint32_t ring = smem_addr(smem_raw) + tid * VB;
  T* s_head = reinterpret_cast<T*>(smem_raw + RG::kBytes);    // [GPC][COLS]
  T* s_tail = s_head + GPC * COLS;                            // [GPC][COLS]
  uint32_t* s_key = reinterpret_cast<uint32_t*>(s_tail + GPC * COLS);  // [win]
  uint32_t* s_pos = s_key + win;                              // [win] original position | kCloses | kDirect
  uint32_t* s_uidx = s_pos + win;                             // [win] kMergeAdd only: output row of each position
  __shared__ uint32_t s_flags[GPC];
  __shared__ int s_a, s_b, s_before, s_after, s_nfix, s_fix_c[2], s_fix_len[2];

  const int cta = blockIdx.x, num_ctas = gridDim.x;
  const int d = P.d;
  const int64_t d_ = P.d;
  pdl_wait();
  pdl_launch_dependents();
  wait_for_flags(P.wait.flags, P.wait.epoch, P.wait.world, P.wait.status);
  const int nom_s = cta * stride, nom_e = min(nom_s + stride, n);
  const int w0 = max(nom_s - 1, 0), w1 = min(nom_e + G.slack + 1, n);
  B200EMB_STAMP(0);
  for (int i = w0 + tid; i < w1; i += kBlock) {
    s_key[i - w0] = skey[i];
    s_pos[i - w0] = perm[i];
    if constexpr (OP == UpdateOp::kMergeAdd) s_uidx[i - w0] = P.uidx[i];
  }
  // kMergeAdd: where a merged row goes is a function of its output index and (with a scatter map) a few small
  // per-destination tables: both are staged here so that closing a segment costs no dependent global load
  constexpr int kSmemShards = 64;
  __shared__ int64_t s_sb[OP == UpdateOp::kMergeAdd ? kSmemShards + 1 : 1];
  __shared__ int64_t s_sm[OP == UpdateOp::kMergeAdd ? 2 * kSmemShards : 1];
  __shared__ T* s_sd[OP == UpdateOp::kMergeAdd ? kSmemShards : 1];
  const bool shard_tables = OP == UpdateOp::kMergeAdd && P.shards > 0 && P.shards <= kSmemShards;
  if constexpr (OP == UpdateOp::kMergeAdd) {
    if (shard_tables) {
      if (tid <= P.shards) s_sb[tid] = P.shard_bounds[tid];
      if (tid < P.shards) {
        s_sd[tid] = P.shard_dst[tid];
        s_sm[2 * tid] = P.shard_meta ? P.shard_meta[2 * tid] : 0;
        s_sm[2 * tid + 1] = P.shard_meta ? P.shard_meta[2 * tid + 1] : (int64_t(1) << 62);
      }
    }
  }
  __syncthreads();
  const auto K = [&](int i) { return s_key[i - w0]; };          // w0 <= i < w1
  // kMergeAdd: destination of the merged row that closes at sorted position idx (nullptr: its owner refused it)
  const auto merge_dst_of = [&](int idx) -> T* {
    const int64_t row = s_uidx[idx - w0];
    if (P.shards == 0) return P.param + row * d_;
    if (!shard_tables) {
      int g = 0;
      while (g + 1 < P.shards && row >= P.shard_bounds[g + 1]) ++g;
      int64_t k = row - P.shard_bounds[g];
      if (P.shard_meta) {
        if (k >= P.shard_meta[2 * g + 1]) return nullptr;
        k += P.shard_meta[2 * g];
      }
      return P.shard_dst[g] + k * d_;
    }
    int g = 0;
    while (g + 1 < P.shards && row >= s_sb[g + 1]) ++g;
    const int64_t k = row - s_sb[g];
    if (k >= s_sm[2 * g + 1]) return nullptr;
    return s_sd[g] + (k + s_sm[2 * g]) * d_;
  };
  // adjusted boundary of nominal position p (the CTAs on both sides of it compute the same value)
  if (warp < 2) {
    const int p = warp == 0 ? nom_s : nom_e;
    int adj = p;
    if (p > 0 && p < n && K(p) == K(p - 1)) {
      const int hi = min(p + G.slack, n - 1);
      for (int q0 = p + 1; q0 <= hi; q0 += 32) {
        const int q = q0 + lane;
        const bool brk = q <= hi && K(q) != K(q - 1);
        const uint32_t m = __ballot_sync(0xffffffffu, brk);
        if (m) { adj = q0 + __ffs(m) - 1; break; }
      }
    }
    if (lane == 0) { if (warp == 0) s_a = adj; else s_b = adj; }
  }
  __syncthreads();
  const int a = s_a, b = s_b;                                   // this CTA reduces sorted positions [a, b)
  const uint32_t prev_key = a > 0 ? K(a - 1) : kNoKey;
  const uint32_t next_key = b < n ? K(b) : kNoKey;

  // ---- chains: segments cut by CTA boundaries, known from the keys alone ------------------------
  // A chain is one segment spread over consecutive CTAs: it LEAVES the CTA where it starts (tail
  // carry) and ENTERS the following ones (head carries).  Every member counts itself in at
  // arrive[first CTA of the chain]; the member that arrives last closes the chain (close_chain).
  // No CTA ever waits for another one.  The chain's extent is found by probing the sorted keys at
  // the CTA boundaries: the first 32 probes are issued here and only read after the main loop.
  const uint32_t k_first = K(a), k_last = K(b - 1);
  // Out-of-range rows (key == height: flagged by the plan, and the padding of the sharded owner side) sort
  // last and are never applied: a CTA that holds nothing else is done, and they never form a chain.
  if (k_first >= P.height) return;
  const bool has_head = prev_key == k_first;                          // a segment enters from CTA cta-1
  const bool through = has_head && k_last == k_first && next_key == k_first;   // ... and runs on past this CTA
  const bool has_tail = !through && next_key == k_last && k_last < P.height;   // a segment that starts here leaves
  // cut(j, k): the boundary before CTA j stayed inside a segment of key k (it was not moved): the keys
  // from position j*stride-1 to `slack` elements further on are all k
  const auto cut = [&](int j, uint32_t k) {
    const int64_t p = static_cast<int64_t>(j) * stride;
    return __ldg(skey + p - 1) == k && __ldg(skey + min(p + G.slack, static_cast<int64_t>(n) - 1)) == k;
  };
  const auto enters = [&](int j) { return j >= 1 && cut(j, k_first); };     // the chain enters CTA j from j-1
  const uint32_t k_fwd = through ? k_first : k_last;
  const auto member = [&](int j) { return j < num_ctas && cut(j, k_fwd); }; // CTA j starts inside the chain
  bool probe = false;
  if (warp == 0 && has_head) probe = enters(cta - 1 - lane);
  if (warp == 1 && (through || has_tail)) probe = member(cta + 1 + lane);

  const int chunk_c = (b - a + GPC - 1) / GPC;                  // this CTA's elements per group
  const int ngroups = (b - a + chunk_c - 1) / chunk_c;          // groups that own elements
  const int gs = a + g * chunk_c;
  const int ge = min(gs + chunk_c, b);
  const bool active = g < ngroups;
  uint32_t prev = kNoKey, next = kNoKey, kf = 0, last = 0;
  bool enters_chunk = false, open_tail = false;
  if (active) {
    prev = gs > a ? K(gs - 1) : prev_key;
    next = ge < b ? K(ge) : next_key;
    kf = K(gs);
    last = K(ge - 1);
    enters_chunk = kf == prev;                                  // the chunk starts inside a segment begun before it
    open_tail = last == next && !(enters_chunk && last == kf);  // a segment starts in this chunk and runs on
    // per-element flags, from the keys alone: where segments end, and whether they began in this chunk
    for (int idx = gs + l; idx < ge; idx += LPR) {
      const uint32_t key = K(idx);
      const uint32_t kn = idx + 1 < ge ? K(idx + 1) : next;
      uint32_t f = 0;
      if (key != kn) f = kCloses | ((enters_chunk && key == kf) ? 0u : kDirect);
      s_pos[idx - w0] |= f;
    }
  }
  __syncwarp();       // a group lives inside one warp
  B200EMB_STAMP(1);
  const int nbatch = active ? (ge - gs + B - 1) / B : 0;

  for (int col0 = 0; col0 < d; col0 += COLS) {
    uint32_t flags = 0, tail_key = 0, tail_eref = 0;
    if (active) {
      // state rows of a key into ring slot `slot` (parts 1..NS)
      const auto fetch_state = [&](uint32_t slot_addr, uint32_t key) {
        if constexpr (NS > 0) {
          if (key >= P.height) return;
          const int64_t base = static_cast<int64_t>(key) * d;
#pragma unroll
          for (int kk = 0; kk < NV; ++kk) {
            const int col = col0 + (l + kk * LPR) * EPV;
            if (col >= d) continue;
            const uint32_t dst = slot_addr + kPartStride + kk * kVecStride;
            cp_async_s<VB>(dst, P.param + base + col);
            if constexpr (OP == UpdateOp::kAdam) {
              cp_async_s<VB>(dst + kPartStride, P.m1 + base + col);
              cp_async_s<VB>(dst + 2 * kPartStride, P.m2 + base + col);
            }
            if constexpr (OP == UpdateOp::kAdagrad) cp_async_s<VB>(dst + kPartStride, P.m1 + base + col);
          }
        }
      };
      const auto read_state = [&](uint32_t slot_addr, uint32_t key, RowState<T, OP, NV, EPV>& st) {
        if constexpr (NS > 0) {
          if (key >= P.height) return;
#pragma unroll
          for (int q = 0; q < NS; ++q)
#pragma unroll
            for (int kk = 0; kk < NV; ++kk) {
              const int col = col0 + (l + kk * LPR) * EPV;
              if (col < d) lds_vec(slot_addr + (1 + q) * kPartStride + kk * kVecStride, st.v[q][kk]);
            }
        }
      };
      // Everything about a batch is known from the staged keys, so the state rows of every segment that
      // closes in it are requested together with its gradient rows, kRingStages batches ahead.
      const auto request = [&](int bi) {
        const int i0 = gs + bi * B;
        const uint32_t stage_addr = ring + (bi % S) * (B * kSlotStride);
#pragma unroll
        for (int q = 0; q < B; ++q) {
          const int idx = i0 + q;
          if (idx < ge) {
            const uint32_t pw = s_pos[idx - w0];
            const T* src = value + static_cast<int64_t>(pw & kPosMask) * d;
            const uint32_t slot_addr = stage_addr + q * kSlotStride;
#pragma unroll
            for (int kk = 0; kk < NV; ++kk) {
              const int col = col0 + (l + kk * LPR) * EPV;
              if (col < d) cp_async_s<VB>(slot_addr + kk * kVecStride, src + col);
            }
            if (RG::kStateInRing && (pw & kDirect)) fetch_state(slot_addr, K(idx));
          }
        }
      };
      if (RG::kTailInRing && open_tail) fetch_state(ring + kTailSlot * kSlotStride, last);   // joins the first group
#pragma unroll
      for (int s = 0; s < S; ++s) {   // empty groups keep the wait arithmetic uniform
        if (s < nbatch) request(s);
        cp_async_commit();
      }
      bool first = true;   // no segment has closed inside this chunk yet
      T acc[NV][EPV];
      zero_acc<T, NV, EPV>(acc);
      for (int bi = 0; bi < nbatch; ++bi) {
        cp_async_wait<S - 1>();       // batch bi has landed (this thread's own copies)
        const int i0 = gs + bi * B;
        const uint32_t stage_addr = ring + (bi % S) * (B * kSlotStride);
#pragma unroll
        for (int q = 0; q < B; ++q) {
          const int idx = i0 + q;
          if (idx < ge) {
            const uint32_t slot_addr = stage_addr + q * kSlotStride;
#pragma unroll
            for (int kk = 0; kk < NV; ++kk) {
              const int col = col0 + (l + kk * LPR) * EPV;
              if (col < d) {
                T r[EPV];
                lds_vec(slot_addr + kk * kVecStride, r);
#pragma unroll
                for (int e = 0; e < EPV; ++e) acc[kk][e] = A::add(acc[kk][e], r[e]);
              }
            }
            const uint32_t pw = s_pos[idx - w0];
            if (pw & kCloses) {   // the segment closes here
              if (pw & kDirect) {
                const uint32_t key = K(idx);
                if constexpr (OP == UpdateOp::kMergeAdd) {
                  RowState<T, OP, NV, EPV> st;
                  apply_state<T, OP, LPR, NV, VEC>(P, key, static_cast<uint32_t>(idx), col0, l, acc, st,
                                                   merge_dst_of(idx), true);
                } else if constexpr (RG::kStateInRing) {
                  RowState<T, OP, NV, EPV> st;
                  read_state(slot_addr, key, st);
                  apply_state<T, OP, LPR, NV, VEC>(P, key, static_cast<uint32_t>(idx), col0, l, acc, st);
                } else {
                  consume<T, OP, LPR, NV, VEC>(P, key, static_cast<uint32_t>(idx), col0, l, acc);
                }
              } else {  // first segment of the chunk, entered from the previous chunk
#pragma unroll
                for (int kk = 0; kk < NV; ++kk)
#pragma unroll
                  for (int e = 0; e < EPV; ++e) s_head[g * COLS + (l + kk * LPR) * EPV + e] = acc[kk][e];
                flags |= kHead;
              }
              zero_acc<T, NV, EPV>(acc);
              first = false;
            }
          }
        }
        if (bi + S < nbatch) request(bi + S);   // refill the stage just drained
        cp_async_commit();
      }
      cp_async_wait<0>();
      B200EMB_STAMP(2);
      if (last == next) {  // the chunk ends inside a segment: leave a partial sum
        const bool left_open = first && last == prev;
        T* dst = (left_open ? s_head : s_tail) + g * COLS;
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int e = 0; e < EPV; ++e) dst[(l + kk * LPR) * EPV + e] = acc[kk][e];
        if (left_open) {
          flags |= kHead | kThrough;
        } else {
          flags |= kTail;
          tail_key = last;
          tail_eref = static_cast<uint32_t>(ge - 1);
        }
      }
    }
    if (l == 0) s_flags[g] = flags;
    __syncthreads();
    B200EMB_STAMP(3);

    // ---- combine partial segments across the CTA's groups, in group order -----------------
    if (active && (flags & kTail)) {
      T sum[NV][EPV];
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int e = 0; e < EPV; ++e) sum[k][e] = s_tail[g * COLS + (l + k * LPR) * EPV + e];
      int j = g + 1;
      bool to_end = true;
      while (j < ngroups) {
        const uint32_t f = s_flags[j];
        if (!(f & kHead)) { to_end = false; break; }
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int e = 0; e < EPV; ++e)
            sum[k][e] = A::add(sum[k][e], s_head[j * COLS + (l + k * LPR) * EPV + e]);
        if (!(f & kThrough)) { to_end = false; break; }
        ++j;
      }
      if (to_end) {  // the segment runs past the CTA's last element: leave a carry
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          if (col < d) store_vec<T, VEC>(C.tail + static_cast<int64_t>(blockIdx.x) * d + col, sum[k]);
        }
      } else if constexpr (RG::kTailInRing) {   // its state rows were requested before the main loop
        RowState<T, OP, NV, EPV> st;
        if (tail_key < P.height) {
#pragma unroll
          for (int q = 0; q < NS; ++q)
#pragma unroll
            for (int kk = 0; kk < NV; ++kk) {
              const int col = col0 + (l + kk * LPR) * EPV;
              if (col < d) lds_vec(ring + kTailSlot * kSlotStride + (1 + q) * kPartStride + kk * kVecStride, st.v[q][kk]);
            }
        }
        apply_state<T, OP, LPR, NV, VEC>(P, tail_key, tail_eref, col0, l, sum, st);
      } else if constexpr (OP == UpdateOp::kMergeAdd) {
        RowState<T, OP, NV, EPV> st;
        apply_state<T, OP, LPR, NV, VEC>(P, tail_key, tail_eref, col0, l, sum, st,
                                         merge_dst_of(static_cast<int>(tail_eref)), true);
      } else {
        consume<T, OP, LPR, NV, VEC>(P, tail_key, tail_eref, col0, l, sum);
      }
    }
    if (g == 0 && active && (flags & kHead)) {  // segment entering from the previous CTA
      T sum[NV][EPV];
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int e = 0; e < EPV; ++e) sum[k][e] = s_head[(l + k * LPR) * EPV + e];
      bool to_end = (flags & kThrough) != 0;
      int j = 1;
      while (to_end && j < ngroups) {
        const uint32_t f = s_flags[j];
        if (!(f & kHead)) { to_end = false; break; }
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int e = 0; e < EPV; ++e)
            sum[k][e] = A::add(sum[k][e], s_head[j * COLS + (l + k * LPR) * EPV + e]);
        if (!(f & kThrough)) { to_end = false; break; }
        ++j;
      }
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int col = col0 + (l + k * LPR) * EPV;
        if (col < d) store_vec<T, VEC>(C.head + static_cast<int64_t>(blockIdx.x) * d + col, sum[k]);
      }
    }
    __syncthreads();
  }
  B200EMB_STAMP(4);
  if (num_ctas == 1) return;

  // ---- arrivals -----------------------------------------------------------------------------------
  if (!has_head && !has_tail) return;    // no segment crosses this CTA's boundaries
  if (warp == 0 && has_head) {           // CTAs of the chain before this one
    const uint32_t stop = __ballot_sync(0xffffffffu, !probe);
    int before = stop ? __ffs(stop) : 0;
    if (!stop)                           // longer than 32: enters(cta - m) is true for m <= 32, false at m = cta
      before = 1 + warp_last_true(32, cta, lane, [&](int m) { return enters(cta - m); });
    if (lane == 0) s_before = before;
  }
  if (warp == 1 && (through || has_tail)) {   // CTAs of the chain after this one
    const uint32_t stop = __ballot_sync(0xffffffffu, !probe);
    int after = stop ? __ffs(stop) - 1 : 0;
    if (!stop)
      after = warp_last_true(32, num_ctas - cta, lane, [&](int m) { return member(cta + m); });
    if (lane == 0) s_after = after;
  }
  __syncthreads();
  if (tid == 0) {
    // Every thread's carry stores happen before the barrier above; this one fence (cumulative) publishes
    // them device-wide before the CTA counts itself in — the grid-sync idiom, one MEMBAR per CTA instead of 256.
    __threadfence();
    // both arrivals are issued before either count is read: one round trip, not two, for a CTA that a segment
    // enters and another one leaves
    const int c0 = cta - s_before, members = s_before + (through ? s_after : 0) + 1;
    uint32_t seen_head = 0, seen_tail = 0;
    if (has_head) seen_head = atomicAdd(C.arrive + c0, 1u);
    if (has_tail) seen_tail = atomicAdd(C.arrive + cta, 1u);
    int nfix = 0;
    if (has_head && seen_head == static_cast<uint32_t>(members - 1)) {
      s_fix_c[nfix] = c0;
      s_fix_len[nfix++] = members - 1;
    }
    if (has_tail && seen_tail == static_cast<uint32_t>(s_after)) {
      s_fix_c[nfix] = cta;
      s_fix_len[nfix++] = s_after;
    }
    // the other members' carry rows were published before their arrivals: this fence, by the thread that observed
    // them, and the barrier below order the closing loads (L2 loads, ld.cg) of every thread after them
    if (nfix) __threadfence();
    s_nfix = nfix;
  }
  __syncthreads();
  const int nfix = s_nfix;
  B200EMB_STAMP(5);
  if (nfix == 0) return;
  for (int f = 0; f < nfix; ++f) {
    const int c0 = s_fix_c[f];
    const bool own_tail = c0 == cta;     // the chain that starts here; otherwise the one that entered
    close_chain<T, OP, LPR, NV, VEC>(c0, s_fix_len[f], own_tail ? k_last : k_first,
                                     static_cast<uint32_t>(own_tail ? b - 1 : a), P, C, s_head);
    if (tid == 0) C.arrive[c0] = 0;      // ready for the next launch on this workspace
  }
  B200EMB_STAMP(6);
}

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__global__ void __launch_bounds__(kBlock, NV == 1 ? 3 : 2)
k_chunk_reduce(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ perm, ChunkGeom G,
               const T* __restrict__ value, Params<T> P, Carry<T> C) {
  chunk_reduce_body<T, OP, LPR, NV, VEC>(skey, perm, G, value, P, C);   // every exit of the body is CTA-uniform
  signal_peers(P.signal.flag_dst, P.signal.payload_dst, P.signal.starts, P.signal.epoch, P.signal.done_ctr,
               P.signal.world);
}

// ------------------------------------------------------------------------------------------
// Bit-exact order mode: the duplicates of a row are added strictly in input order, so the adds of one element
// form one dependent chain — but the LOADS do not, and neither do different elements.  Rows with up to kLongSeg
// duplicates are walked by one lane group each (k_segment_det); a group that finds more than that drops the row
// (nothing written yet) into a list for k_long_segment_det below.  Same arithmetic, same order, same bits.
constexpr int kLongSeg = 64;

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void det_begin(const Params<T>& P, uint32_t key, int col0, int l,
                                          T (&acc)[NV][Epv<T, VEC>::value]) {
  constexpr int EPV = Epv<T, VEC>::value;
  if constexpr (OP == UpdateOp::kSgd || OP == UpdateOp::kAddToTensor) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int col = col0 + (l + k * LPR) * EPV;
      if (col < P.d) load_vec<T, VEC>(P.param + static_cast<int64_t>(key) * P.d + col, acc[k]);
      else {
#pragma unroll
        for (int e = 0; e < EPV; ++e) acc[k][e] = T(0);
      }
    }
  } else {
    zero_acc<T, NV, EPV>(acc);  // MergeAdd starts from +0.0 (SetConstant, selected_rows_functor.cc:211-212)
  }
}

template <typename T, UpdateOp OP, int NV, int EPV>
__device__ __forceinline__ void det_add(T (&acc)[NV][EPV], const T (&r)[NV][EPV], T lr) {
  using A = Arith<T>;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int e = 0; e < EPV; ++e) {
      if constexpr (OP == UpdateOp::kSgd) acc[k][e] = A::sub(acc[k][e], A::mul(lr, r[k][e]));  // sgd_op.h:77-79
      else acc[k][e] = A::add(acc[k][e], r[k][e]);
    }
}

// the same step with the product lr * g already formed (by whichever thread fetched the row: same operands, same bits)
template <typename T, UpdateOp OP, int NV, int EPV>
__device__ __forceinline__ void det_add_scaled(T (&acc)[NV][EPV], const T (&r)[NV][EPV]) {
  using A = Arith<T>;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int e = 0; e < EPV; ++e) {
      if constexpr (OP == UpdateOp::kSgd) acc[k][e] = A::sub(acc[k][e], r[k][e]);
      else acc[k][e] = A::add(acc[k][e], r[k][e]);
    }
}

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__device__ __forceinline__ void det_end(const Params<T>& P, uint32_t key, uint32_t first_pos, int col0, int l,
                                        T (&acc)[NV][Epv<T, VEC>::value]) {
  constexpr int EPV = Epv<T, VEC>::value;
  if constexpr (OP == UpdateOp::kSgd || OP == UpdateOp::kAddToTensor) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int col = col0 + (l + k * LPR) * EPV;
      if (col < P.d) store_vec<T, VEC>(P.param + static_cast<int64_t>(key) * P.d + col, acc[k]);
    }
  } else {
    consume<T, OP, LPR, NV, VEC>(P, key, first_pos, col0, l, acc);
  }
}

template <typename T, UpdateOp OP, int LPR, int NV, bool VEC>
__global__ void __launch_bounds__(kBlock)
k_segment_det(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ perm, int n,
              const uint32_t* __restrict__ seg_start, const PlanHeader* hdr,
              const T* __restrict__ value, Params<T> P, uint32_t* __restrict__ long_list,
              uint32_t* __restrict__ long_count) {
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int GPC = kBlock / LPR;
  constexpr int COLS = LPR * NV * EPV;
  constexpr int kPrefetch = NV >= 4 ? 2 : (NV == 2 ? 4 : 8);
  const int l = threadIdx.x % LPR;
  const int d = P.d;
  wait_for_flags(P.wait.flags, P.wait.epoch, P.wait.world, P.wait.status);
  const int n_unique = static_cast<int>(hdr->n_unique);
  for (int s = blockIdx.x * GPC + threadIdx.x / LPR; s < n_unique; s += gridDim.x * GPC) {
    const int b = static_cast<int>(seg_start[s]);
    const uint32_t key = skey[b];
    bool is_long = false;
    for (int col0 = 0; col0 < d && !is_long; col0 += COLS) {
      T acc[NV][EPV];
      det_begin<T, OP, LPR, NV, VEC>(P, key, col0, l, acc);
      T lr = T(0);
      if constexpr (OP == UpdateOp::kSgd) lr = *P.lr;
      bool more = true;
      for (int e0 = b; more; e0 += kPrefetch) {
        if (e0 - b >= kLongSeg) {   // col0 == 0 here: a long row leaves on its first column slice
          is_long = true;
          break;
        }
        T r[kPrefetch][NV][EPV];
        bool same[kPrefetch];
#pragma unroll
        for (int q = 0; q < kPrefetch; ++q) {
          const int idx = e0 + q;
          same[q] = idx < n && skey[idx] == key;
          const int64_t src = same[q] ? static_cast<int64_t>(perm[idx]) * d : 0;
#pragma unroll
          for (int k = 0; k < NV; ++k) {
            const int col = col0 + (l + k * LPR) * EPV;
            if (same[q] && col < d) load_vec_stream<T, VEC>(value + src + col, r[q][k]);
            else {
#pragma unroll
              for (int e = 0; e < EPV; ++e) r[q][k][e] = T(0);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < kPrefetch; ++q) {
          if (more && same[q]) det_add<T, OP, NV, EPV>(acc, r[q], lr);
          else more = false;
        }
      }
      if (!is_long) det_end<T, OP, LPR, NV, VEC>(P, key, static_cast<uint32_t>(b), col0, l, acc);
    }
    if (is_long && l == 0) long_list[atomicAdd(long_count, 1u)] = static_cast<uint32_t>(s);
  }
}

// Long rows.  The adds of one ELEMENT of a row are a dependent chain, but the elements are independent, and one SM
// pulls scattered rows out of HBM at a small fraction of the device bandwidth (its loads in flight divided by
// the latency).  So a listed row is cut into column slices of kSliceLanes vectors (64 bytes for fp32), and
// each (row, slice) pair is one work item with a CTA to itself: a hot row is fetched by several SMs at once,
// and a CTA's 64 KB tile holds 1024 duplicates of a slice instead of 256 whole rows.  Thread (rr, l) fetches
// duplicates rr, rr + RPP, ... of the tile into registers while the slice's first lanes add the previous tile
// out of shared memory — in sorted (= input) order, which is all that bit-exactness needs; the sorted positions
// of a tile are read one tile further ahead, so each tile costs one memory round trip, not two.
// ctrs[0] = list length (reset by the last CTA to finish, ctrs[1] counts finished CTAs): both are zero again
// when the kernel ends.
template <bool VEC> struct LongDet {
  static constexpr int kSliceLanes = VEC ? 4 : 16;   // lanes (vectors) per slice
  static constexpr int kJ = 16;                      // duplicates in flight per thread
};
constexpr int kLongDetSmem = 64 * 1024;

template <typename T, UpdateOp OP, bool VEC>
__global__ void __launch_bounds__(kBlock, 1)
k_long_segment_det(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ perm, int n,
                   const uint32_t* __restrict__ seg_start, const T* __restrict__ value, Params<T> P,
                   const uint32_t* __restrict__ long_list, uint32_t* __restrict__ ctrs) {
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int SL = LongDet<VEC>::kSliceLanes;
  constexpr int SC = SL * EPV;                     // columns per slice
  constexpr int RPP = kBlock / SL;                 // duplicates per pass of the CTA
  constexpr int J = LongDet<VEC>::kJ;
  constexpr int R = RPP * J;                       // duplicates per tile
  static_assert(R * SC * sizeof(T) <= kLongDetSmem, "tile fits the dynamic shared memory");
  extern __shared__ __align__(16) unsigned char det_smem[];
  T (*s_tile)[SC] = reinterpret_cast<T (*)[SC]>(det_smem);
  __shared__ uint32_t s_cnt[2];
  const int tid = threadIdx.x, rr = tid / SL, l = tid % SL;
  const int d = P.d;
  const uint32_t count = *reinterpret_cast<volatile uint32_t*>(ctrs);
  const uint32_t slices = static_cast<uint32_t>((d + SC - 1) / SC);
  T lr = T(0);
  if constexpr (OP == UpdateOp::kSgd) lr = *P.lr;
  for (uint32_t item = blockIdx.x; item < count * slices; item += gridDim.x) {
    const int b = static_cast<int>(seg_start[long_list[item / slices]]);
    const int col0 = static_cast<int>(item % slices) * SC;
    const int col = col0 + l * EPV;                // this thread's columns
    const uint32_t key = skey[b];
    T acc[1][EPV];
    if (rr == 0) det_begin<T, OP, SL, 1, VEC>(P, key, col0, l, acc);
    T r[J][EPV];
    uint32_t pos[J];                               // original positions of a tile's duplicates, kNoKey past the end
    bool same[J];
    auto fetch_pos = [&](int base) {
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const int idx = base + j * RPP + rr;
        const bool in = idx < n;
        const uint32_t k2 = in ? skey[idx] : ~key;
        const uint32_t p2 = in ? perm[idx] : 0u;
        pos[j] = k2 == key ? p2 : kNoKey;
      }
    };
    auto fetch = [&]() {
#pragma unroll
      for (int j = 0; j < J; ++j) {
        same[j] = pos[j] != kNoKey;
        if (same[j] && col < d) load_vec_stream<T, VEC>(value + static_cast<int64_t>(pos[j]) * d + col, r[j]);
        else {
#pragma unroll
          for (int e = 0; e < EPV; ++e) r[j][e] = T(0);
        }
      }
    };
    fetch_pos(b);
    fetch();
    fetch_pos(b + R);
    if (tid < 2) s_cnt[tid] = 0u;
    __syncthreads();
    int valid = R;                                 // duplicates in the tile being added; CTA-uniform
    for (int t = 0; valid == R; ++t) {
      uint32_t mine = 0;
#pragma unroll
      for (int j = 0; j < J; ++j) {
        if constexpr (OP == UpdateOp::kSgd) {      // sgd_op.h:77-79: param -= lr * g, the product formed here
#pragma unroll
          for (int e = 0; e < EPV; ++e) r[j][e] = Arith<T>::mul(lr, r[j][e]);
        }
        store_vec<T, VEC>(&s_tile[j * RPP + rr][l * EPV], r[j]);
        mine += same[j] ? 1u : 0u;
      }
      if (l == 0 && mine) atomicAdd(&s_cnt[t & 1], mine);
      if (tid == 0) s_cnt[(t + 1) & 1] = 0u;       // last read before the previous tile's closing barrier
      __syncthreads();
      valid = static_cast<int>(s_cnt[t & 1]);      // a sorted run: the valid duplicates are rows [0, valid)
      if (valid == R) {                            // a full tile: the row may go on
        fetch();
        fetch_pos(b + (t + 2) * R);
      }
      if (rr == 0) {
        // the chain: a block of duplicates is read from shared memory back to back (ordered asm loads), then added
        constexpr int KB = 16;
        constexpr uint32_t kRowBytes = SC * sizeof(T);
        const uint32_t mine_addr = smem_addr(det_smem) + l * EPV * static_cast<uint32_t>(sizeof(T));
        int row = 0;
        for (; row + KB <= valid; row += KB) {
          T v[KB][1][EPV];
#pragma unroll
          for (int q = 0; q < KB; ++q) lds_vec(mine_addr + (row + q) * kRowBytes, v[q][0]);
#pragma unroll
          for (int q = 0; q < KB; ++q) det_add_scaled<T, OP, 1, EPV>(acc, v[q]);
        }
        for (; row < valid; ++row) {
          T v[1][EPV];
          lds_vec(mine_addr + row * kRowBytes, v[0]);
          det_add_scaled<T, OP, 1, EPV>(acc, v);
        }
      }
      __syncthreads();
    }
    if (rr == 0) det_end<T, OP, SL, 1, VEC>(P, key, static_cast<uint32_t>(b), col0, l, acc);
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(ctrs + 1, 1u) == gridDim.x - 1) {
      ctrs[0] = 0u;
      ctrs[1] = 0u;
    }
  }
}

template <typename T, UpdateOp OP>
cudaError_t launch_op(const SortedPlan& plan, const T* value, int64_t n, int64_t d, int64_t height,
                      const UpdateArgs<T>& a, bool deterministic, char* ws, const Workspace& L,
                      cudaStream_t s) {
  Params<T> P{a.param, a.moment1, a.moment2, a.lr, a.beta1_pow, a.beta2_pow,
              a.beta1, a.beta2, a.epsilon, a.uidx, static_cast<int>(d),
              static_cast<uint32_t>(height), a.shard_bounds, a.shard_dst, a.shards, a.shard_meta, a.wait, a.signal};
  Carry<T> C{reinterpret_cast<uint32_t*>(ws + L.off_cta_arrive),
             reinterpret_cast<T*>(ws + L.off_cta_head), reinterpret_cast<T*>(ws + L.off_cta_tail)};
  bool aligned = is_aligned16(value) && (a.shards > 0 || is_aligned16(a.param)) && a.dst_aligned16;
  if (OP == UpdateOp::kAdam) aligned = aligned && is_aligned16(a.moment1) && is_aligned16(a.moment2);
  if (OP == UpdateOp::kAdagrad) aligned = aligned && is_aligned16(a.moment1);
  cudaError_t err = cudaSuccess;
  const int ni = static_cast<int>(n);
  dispatch_row_shape<T>(d, aligned, [&]<int LPR, int NV, bool VEC>() {
    constexpr int GPC = kBlock / LPR;
    constexpr int COLS = LPR * NV * Epv<T, VEC>::value;
    if (deterministic) {
      const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, GPC), 148 * 8));
      // scratch of the fast path, idle in this mode: arrival words 0/1 (zero between launches) count the
      // listed rows and the finished CTAs, the head-carry block holds the list (at most n / kLongSeg
      // entries; the block has max_ctas * d elements >= n / 32 words)
      uint32_t* ctrs = C.arrive;
      uint32_t* list = reinterpret_cast<uint32_t*>(C.head);
      const uint32_t* seg_start = reinterpret_cast<const uint32_t*>(ws + L.off_seg_start);
      k_segment_det<T, OP, LPR, NV, VEC><<<grid, kBlock, 0, s>>>(
          plan.keys, plan.perm, ni, seg_start, reinterpret_cast<const PlanHeader*>(ws + L.off_plan), value, P,
          list, ctrs);
      if (n >= kLongSeg) {   // a run of exactly kLongSeg duplicates is already handed over
        auto long_kern = k_long_segment_det<T, OP, VEC>;
        static std::atomic<unsigned long long> long_attr_done{0};   // > 48 KB of dynamic shared memory: once per device
        int dev = 0;
        cudaGetDevice(&dev);
        const unsigned long long bit = 1ull << (dev & 63);
        if (!(long_attr_done.load(std::memory_order_relaxed) & bit)) {
          if (cudaFuncSetAttribute(long_kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kLongDetSmem) !=
              cudaSuccess) {
            err = cudaGetLastError();
            return;
          }
          long_attr_done.fetch_or(bit, std::memory_order_relaxed);
        }
        long_kern<<<148 * 2, kBlock, kLongDetSmem, s>>>(plan.keys, plan.perm, ni, seg_start, value, P, list, ctrs);
      }
    } else {
      // one wave of lane groups at 3 CTAs per SM.  The chunk size fixes the summation tree, so it must
      // depend on (N, D) alone — not on the consumer (the fast-mode optimizers must reduce in exactly
      // MergeAdd's order); how many CTAs of a given consumer really fit an SM only changes the timing.
      static const int chunk_override = [] {
        const char* ov = std::getenv("B200EMB_CHUNK");   // tuning override, read once
        const int c = ov ? std::atoi(ov) : 0;
        return (c >= kMinChunk && c <= kMaxChunk) ? c : 0;
      }();
      static const int cps = [] {   // CTAs per SM the one-wave geometry aims at (tuning override, read once)
        const char* ov = std::getenv("B200EMB_REDUCE_CPS");
        const int c = ov ? std::atoi(ov) : 0;
        return (c >= 1 && c <= 4) ? c : 3;
      }();
      const int chunk = chunk_override ? chunk_override : choose_chunk(n, GPC, cps);
      const int ctas = static_cast<int>(ceil_div(n, static_cast<int64_t>(GPC) * chunk));
      const int slack = std::min(kBoundarySlack, GPC * chunk / 2);
      using RG = Ring<T, OP, NV, VEC>;
      const size_t smem = RG::kBytes + sizeof(T) * 2 * GPC * COLS +
                          sizeof(uint32_t) * (OP == UpdateOp::kMergeAdd ? 3 : 2) * (GPC * chunk + slack + 2);
      const ChunkGeom G{ni, chunk, slack};
      auto kern = k_chunk_reduce<T, OP, LPR, NV, VEC>;
      // opt in to > 48 KB of dynamic shared memory once per device (the attribute is per device)
      static std::atomic<unsigned long long> attr_done{0};
      int dev = 0;
      cudaGetDevice(&dev);
      const unsigned long long bit = 1ull << (dev & 63);
      if (!(attr_done.load(std::memory_order_relaxed) & bit)) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) {
          err = cudaGetLastError();
          return;
        }
        attr_done.fetch_or(bit, std::memory_order_relaxed);
      }
      err = launch_row_kernel(kern, ctas, kBlock, smem, s, plan.keys, plan.perm, G, value, P, C);
      return;
    }
    err = cudaGetLastError();
  });
  return err;
}

}  // namespace

template <typename T>
cudaError_t launch_update(const SortedPlan& plan, const T* value, int64_t n, int64_t d,
                          int64_t height, const UpdateArgs<T>& a, bool deterministic, char* ws,
                          const Workspace& L, int32_t* /*status*/, cudaStream_t s) {
  switch (a.op) {
    case UpdateOp::kMergeAdd:
      return launch_op<T, UpdateOp::kMergeAdd>(plan, value, n, d, height, a, deterministic, ws, L, s);
    case UpdateOp::kAddToTensor:
      return launch_op<T, UpdateOp::kAddToTensor>(plan, value, n, d, height, a, deterministic, ws, L, s);
    case UpdateOp::kSgd:
      return launch_op<T, UpdateOp::kSgd>(plan, value, n, d, height, a, deterministic, ws, L, s);
    case UpdateOp::kAdam:
      return launch_op<T, UpdateOp::kAdam>(plan, value, n, d, height, a, deterministic, ws, L, s);
    case UpdateOp::kAdagrad:
      return launch_op<T, UpdateOp::kAdagrad>(plan, value, n, d, height, a, deterministic, ws, L, s);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b200emb
