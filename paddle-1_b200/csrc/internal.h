// internal.h — host-side contracts between capi.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace b200emb {

// ---- radix sort geometry (sort.cu) -----------------------------------------------------
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;
constexpr int kSortTile = kSortThreads * kSortItems;   // 2048 keys per tile
constexpr int kMaxPasses = 4;                          // keys < 2^32
constexpr int kScanTile = 2048;                        // head-flag scan tile

// ---- segmented-reduce geometry (segreduce.cu) ------------------------------------------
constexpr int kMinChunk = 4;     // sorted elements walked by one lane group, lower bound
constexpr int kMaxChunk = 128;
constexpr int kMinGroupsPerCta = 8;   // 256 threads / 32 lanes

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Everything the kernels keep in the caller's workspace.  Offsets in bytes.
struct Workspace {
  // sort scratch (nothing here needs clearing between plans)
  size_t off_digit_counts;  // u32 [passes][sort_tiles + 1][bins]  per-tile digit counts / prefixes
  size_t off_scan_state;    // u32 [scan_tiles]                    head counts per scan tile
  // plan
  size_t off_plan;        // PlanHeader
  size_t off_keys[2];     // u32 [n]
  size_t off_perm[2];     // u32 [n]
  size_t off_uidx;        // u32 [n]   unique index of each sorted position
  size_t off_seg_start;   // u32 [n+1] first sorted position of each unique row
  // carries of the chunked reduce
  size_t off_cta_arrive;  // u32 [max_ctas]  chain arrival counters: cleared by every plan, left at zero by every reduce
  size_t off_cta_head;    // T [max_ctas][d]
  size_t off_cta_tail;    // T [max_ctas][d]
  size_t total;
  int64_t sort_tiles, scan_tiles, max_ctas;
};

struct PlanHeader {      // lives in the workspace (device)
  unsigned long long n_unique;  // written by the head scan
};

Workspace make_workspace(int64_t n, int64_t d, int64_t elem_size);

struct SortedPlan {      // device pointers into the workspace after plan_rows
  const uint32_t* keys;  // sorted ids (invalid ids -> `height`)
  const uint32_t* perm;  // original position of each sorted element (stable)
  int passes;
};

// Optional id -> sort-key remapping applied while narrowing ids to 32 bits.  With `shards` > 1 the
// key is (id % shards) * shard_height + id / shards: the owner convention of the reference's
// split_ids op (operators/split_ids_op.h:45-49), so one sort groups ids by owning shard and, inside
// a shard, by local row.  `raw_height` bounds the incoming ids; the plan's height is then
// shards * shard_height.
struct KeyMap {
  uint32_t shards = 1, shard_height = 0, raw_height = 0;
  bool owner_only = false;   // key = id % shards: a stable partition by owner (split_ids on a LoDTensor)
#if defined(__CUDACC__)
  __host__ __device__
#endif
  uint32_t operator()(uint32_t id) const {
    if (owner_only) return id % shards;
    return shards > 1 ? (id % shards) * shard_height + id / shards : id;
  }
};

// sort.cu
cudaError_t launch_plan(const int64_t* rows, int64_t n, int64_t height, char* ws,
                        const Workspace& L, int32_t* status, cudaStream_t s, SortedPlan* out,
                        const KeyMap* keymap = nullptr);
SortedPlan plan_view(int64_t n, int64_t height, char* ws, const Workspace& L);
cudaError_t launch_head_scan(const SortedPlan& plan, int64_t n, uint32_t invalid_key, char* ws,
                             const Workspace& L, int64_t* out_rows, int64_t* n_unique_out,
                             cudaStream_t s);

// regularize.cu
template <typename T>
cudaError_t launch_sparse_decay(const T* table, int64_t height, int64_t d, const int64_t* rows, int64_t n,
                                T coeff, int mode, const T* addend, T* out, int32_t* status,
                                cudaStream_t s);
int max_sections();
cudaError_t launch_section_of(const int64_t* rows, int64_t n, const int64_t* height_sections, int n_sections,
                              int64_t* out, cudaStream_t s);
cudaError_t launch_rebase_rows(const int64_t* rows, const SortedPlan& plan, int64_t n,
                               const int64_t* height_sections, int n_sections, int64_t* out,
                               cudaStream_t s);

// cluster_sort.cu
bool cluster_sort_fits(int64_t n);
cudaError_t launch_cluster_sort(const int64_t* rows, int64_t n, int64_t height, int passes,
                                uint32_t* keys_out, uint32_t* perm_out, int32_t* status,
                                cudaStream_t s);

// Optional wait at the start of a consumer kernel of the sharded path: the first `world` threads of every CTA
// spin (ld.acquire.sys) until flags[r] >= *epoch for every source rank r — the data those ranks stored into
// this GPU's buffers before raising their flag is then visible to the whole CTA.  flags == nullptr: no wait.
struct WaitSpec {
  const unsigned long long* flags = nullptr;   // [world], written by the peers (st.release.sys)
  const unsigned long long* epoch = nullptr;   // device word: the step being waited for
  int world = 0;
  int32_t* status = nullptr;                   // B200EMB_STATUS_BARRIER_TIMEOUT after ~20 s
};

// Optional publication at the end of a producer kernel of the sharded path.  Every CTA orders its stores to
// the other GPUs before its arrival (barrier, then one cumulative gpu-scope fence); the last CTA to finish writes, per destination
// rank r, the payload {starts[r], starts[r+1]-starts[r]} (if payload_dst) and then the flag = *epoch
// (st.release.sys, the kernel's only system-scope fence, cumulative over all arrivals it observed) — data and
// flag leave from the same kernel.  flag_dst == nullptr: nothing is published.
struct SignalSpec {
  unsigned long long* const* flag_dst = nullptr;   // [world] device table: this rank's flag word on every rank
  int64_t* const* payload_dst = nullptr;           // [world] optional: two int64 per destination
  const int64_t* starts = nullptr;                 // [world + 1] with payload_dst
  const unsigned long long* epoch = nullptr;
  unsigned int* done_ctr = nullptr;                // zero between launches
  int world = 0;
};

// gather.cu
// `seg_start`/`seg_dst` (device arrays, `segs` entries + 1 / `segs` entries) optionally scatter the output:
// row k of segment g (seg_start[g] <= k < seg_start[g+1]) is stored at seg_dst[g] + (k - seg_start[g]) * d.
template <typename T>
cudaError_t launch_gather(const T* table, int64_t height, int64_t d, const int64_t* ids, int64_t n,
                          int64_t padding_idx, T* out, int32_t* status, cudaStream_t s,
                          const int64_t* seg_start = nullptr, T* const* seg_dst = nullptr, int segs = 0,
                          const SignalSpec& signal = SignalSpec());
// with a scatter map only the first seg_start[segs] ids are looked up (a count known on the device alone)

// segreduce.cu
enum class UpdateOp { kMergeAdd, kAddToTensor, kSgd, kAdam, kAdagrad };
template <typename T>
struct UpdateArgs {
  UpdateOp op;
  T* param = nullptr;        // kSgd/kAdam: Param; kAddToTensor: the dense tensor; kMergeAdd: out_value
  T* moment1 = nullptr;
  T* moment2 = nullptr;
  const T* lr = nullptr;
  const T* beta1_pow = nullptr;
  const T* beta2_pow = nullptr;
  T beta1 = 0, beta2 = 0, epsilon = 0;
  const uint32_t* uidx = nullptr;  // kMergeAdd: sorted position -> output row
  const int64_t* shard_bounds = nullptr;   // kMergeAdd scatter map (device arrays), see segreduce.cuh Params
  T* const* shard_dst = nullptr;
  const int64_t* shard_meta = nullptr;     // optional device pairs {row offset inside the destination, rows it accepts}
  int shards = 0;
  WaitSpec wait;                           // sharded path: the gradient rows come from other GPUs
  SignalSpec signal;                       // sharded path: the merged rows go to other GPUs
  bool dst_aligned16 = true;               // caller's promise about the scatter destinations
};
template <typename T>
cudaError_t launch_update(const SortedPlan& plan, const T* value, int64_t n, int64_t d,
                          int64_t height, const UpdateArgs<T>& args, bool deterministic, char* ws,
                          const Workspace& L, int32_t* status, cudaStream_t s);

int choose_chunk(int64_t n, int groups_per_cta, int ctas_per_sm);

// shard.cu
cudaError_t launch_inverse_map(const SortedPlan& plan, const uint32_t* uidx, int64_t n, uint32_t* inv,
                               cudaStream_t s);
cudaError_t launch_shard_split(const int64_t* ukeys, const SortedPlan& plan, char* ws,
                               const Workspace& L, int shards, uint32_t shard_height, int64_t n,
                               uint32_t invalid_key, int64_t* bounds, int64_t* elem_bounds,
                               int64_t* local_rows, cudaStream_t s);
cudaError_t launch_segment_scatter_i64(const int64_t* src, int64_t n_upper, const int64_t* bounds,
                                       int64_t* const* dst, int segs, cudaStream_t s);
// mode 0: out[e] = src[perm[e]]   1: out[perm[e]] = src[uidx[e]]   2: out[perm[e]] = src[e]
template <typename T>
cudaError_t launch_permute_rows(const T* src, const SortedPlan& plan, const uint32_t* uidx, int64_t n,
                                int64_t d, T* out, int mode, cudaStream_t s, const WaitSpec& wait = WaitSpec());

}  // namespace b200emb
