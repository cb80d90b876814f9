/* b200emb.h — C ABI of the B200 (sm_100a) sparse-embedding kernels.
 *
 * Drop-in boundary for PaddlePaddle Fluid's sparse-embedding hot path: each entry point is
 * what the body of one reference `OpKernel<T>::Compute` (or one math functor) would call.
 * Citations are relative to /root/reference/paddle/fluid/.
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer on the current device unless named `*_host`;
 *   - `stream` is a cudaStream_t passed as void*; work is only enqueued, never synchronised;
 *   - nothing is allocated: temporaries live in the caller-provided `workspace`
 *     (size from b200emb_workspace_bytes, 256-byte aligned; the framework hands it out from
 *     its own allocator the way MergeAdd's temporary SelectedRows is, selected_rows_functor.h:63-68);
 *   - return value: 0 on success, a negative B200EMB_E_* code otherwise (never throws);
 *   - ids / rows are int64 (lookup_table_op.h:44); tables are float or double (`_f32` / `_f64`);
 *   - limits: height < 2^32-1, N < 2^30 (the reference caps at 2^31 via int casts,
 *     platform/for_range.h:61).
 *   - `status` (nullable int32*): device word the kernels OR error bits into instead of the
 *     reference's PADDLE_ENFORCE / device assert: bit0 = lookup id outside [0,height)
 *     (lookup_table_op.h:59-60), bit1 = gradient row outside [0,height) (sgd_op.h:75-76).
 *     Offending rows are skipped (lookup writes zeros).
 *
 * flags:
 *   B200EMB_DETERMINISTIC  bit-exact order mode: every duplicate is accumulated in input order
 *                          exactly like the reference CPU loops (MergeAdd: from +0.0,
 *                          selected_rows_functor.cc:214-225; SGD: one p-=lr*g per occurrence,
 *                          sgd_op.h:74-80).  Default (fast) mode reduces duplicates with a
 *                          fixed-shape tree (run-to-run deterministic, differs in rounding).
 *   B200EMB_PLAN_READY     the workspace already holds the sort plan of exactly these
 *                          rows (b200emb_plan_rows was run on it): skip the sort.
 */
#ifndef B200EMB_H_
#define B200EMB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200EMB_OK 0
#define B200EMB_E_INVALID (-1)     /* bad argument (null pointer, negative size, D<=0 ...) */
#define B200EMB_E_WORKSPACE (-2)   /* workspace missing, misaligned or too small */
#define B200EMB_E_CUDA (-3)        /* a CUDA call / launch failed (see b200emb_last_cuda_error) */
#define B200EMB_E_UNSUPPORTED (-4) /* size beyond the documented limits */
#define B200EMB_E_SCRATCH_LIMIT (-5) /* b200emb_sparse_sgd_noalloc_f32: request needs more than the built-in scratch */

#define B200EMB_DETERMINISTIC 1u
#define B200EMB_PLAN_READY 2u
#define B200EMB_UNIQUE_READY 4u   /* with PLAN_READY: the workspace also holds the unique index of
                                     b200emb_plan_unique_rows for these rows: skip the head scan
                                     (out_rows / n_unique of merge_add are then not written) */

#define B200EMB_STATUS_BAD_LOOKUP_ID 1
#define B200EMB_STATUS_BAD_GRAD_ROW 2
#define B200EMB_STATUS_BARRIER_TIMEOUT 4   /* a peer barrier / arrival flag gave up waiting for a rank (~20 s) */
#define B200EMB_STATUS_SHARD_OVERFLOW 8    /* b200emb_shard_*: more rows than the capacities given at create() */

int b200emb_abi_version(void);
/* cudaError_t of the last failing CUDA call made by this library on the calling thread. */
int b200emb_last_cuda_error(void);

/* Bytes of workspace the sort/merge/update entry points need for n rows of width d
 * (elem_size 4 or 8).  Pure host arithmetic. */
size_t b200emb_workspace_bytes(int64_t n, int64_t d, int64_t elem_size);

/* lookup_table forward — replaces LookupTableCUDAKernel::Compute + LookupTable<>
 * (operators/lookup_table_op.cu:24-50,72-103; CPU semantics lookup_table_op.h:36-84):
 * out[k,:] = (ids[k]==padding_idx) ? 0 : table[ids[k],:].  padding_idx = -1 means none. */
int b200emb_lookup_table_f32(const float* table, int64_t height, int64_t d, const int64_t* ids,
                             int64_t n, int64_t padding_idx, float* out, int32_t* status,
                             void* stream);
int b200emb_lookup_table_f64(const double* table, int64_t height, int64_t d, const int64_t* ids,
                             int64_t n, int64_t padding_idx, double* out, int32_t* status,
                             void* stream);

/* lookup_table_grad, sparse branch — replaces LookupTableGradCUDAKernel's is_sparse path
 * (operators/lookup_table_op.cu:114-147; lookup_table_op.h:106-134): rows = ids verbatim,
 * value = copy of dOut.  Either output may alias its input (then that copy is skipped). */
int b200emb_lookup_table_grad_sparse_f32(const int64_t* ids, const float* d_out, int64_t n,
                                         int64_t d, int64_t* rows, float* value, void* stream);
int b200emb_lookup_table_grad_sparse_f64(const int64_t* ids, const double* d_out, int64_t n,
                                         int64_t d, int64_t* rows, double* value, void* stream);

/* `sum` over SelectedRows / SelectedRowsAddTo — replaces operators/sum_op.h:71-145 +
 * math/selected_rows_functor.cu:142-174: appends n_in rows at row offset `row_offset`
 * of (out_rows, out_value).  elem_size = sizeof(T). */
int b200emb_selected_rows_add_to(const int64_t* in_rows, const void* in_value, int64_t n_in,
                                 int64_t d, int64_t elem_size, int64_t* out_rows,
                                 void* out_value, int64_t row_offset, void* stream);

/* Sort plan: stable sort-by-id of `rows` into the workspace (sorted keys + original
 * positions).  Lets the caller overlap the sort with other work (e.g. the forward gather
 * on another stream) and then run merge_add / sparse_sgd / sparse_adam with
 * B200EMB_PLAN_READY.  No counterpart in the reference (its MergeAdd builds a std::set on the
 * host: math/selected_rows_functor.cu:265-267). */
int b200emb_plan_rows(const int64_t* rows, int64_t n, int64_t height, void* workspace,
                      size_t workspace_bytes, int32_t* status, void* stream);

/* Sharded lookup plan (K6) — the device-side counterpart of the reference's CPU-only split_ids op
 * (operators/split_ids_op.h:36-58, `PADDLE_THROW` on GPU at :30-32): owner(id) = id % num_shards,
 * local row = id / num_shards.  Sorts rows by key = owner * ceil(height/num_shards) + local row and
 * de-duplicates them:
 *   unique_keys[0..U)  (nullable unless one of the three outputs below is wanted) ascending unique
 *                      keys (must hold n entries), *n_unique = U (nullable device int64),
 *   shard_bounds[g]    (nullable, num_shards+1 entries) first unique index owned by shard g,
 *   elem_bounds[g]     (nullable, num_shards+1 entries) first position, in plan order, owned by shard g,
 *   local_rows[0..U)   (nullable) key % shard_height = the row inside the owner's shard.
 * The workspace then holds the plan (keys, positions, unique index): run merge_add / sparse_* on the
 * same rows with B200EMB_PLAN_READY and height = plan_height = num_shards*ceil(height/num_shards)
 * (= height when num_shards == 1) to reduce in exactly this order. */
int b200emb_plan_unique_rows(const int64_t* rows, int64_t n, int64_t height, int64_t num_shards,
                             int64_t* unique_keys, int64_t* n_unique, int64_t* shard_bounds,
                             int64_t* elem_bounds, int64_t* local_rows, void* workspace,
                             size_t workspace_bytes, int32_t* status, void* stream);

/* Inverse of the de-duplication — the device-side merge_ids (operators/merge_ids_op.h:63-87,
 * CPU-only in the reference): out[i,:] = unique_values[u(i),:] where u(i) is the unique index of
 * rows[i] in the plan held by `workspace` (b200emb_plan_unique_rows). */
int b200emb_expand_unique_f32(const float* unique_values, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, float* out, void* stream);
int b200emb_expand_unique_f64(const double* unique_values, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, double* out, void* stream);

/* out[e,:] = value[position of the e-th id in plan order,:] — rows of a [n,d] block reordered into
 * the plan's (owner, local row, original position) order; used to ship un-merged gradient rows
 * in the bit-exact mode (split_ids on a SelectedRows, operators/split_ids_op.h:59-92). */
int b200emb_gather_sorted_f32(const float* value, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, float* out, void* stream);
int b200emb_gather_sorted_f64(const double* value, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, double* out, void* stream);

/* split_ids / merge_ids with the reference's exact semantics (both are CPU-only in the
 * reference and PADDLE_THROW on GPU: split_ids_op.h:30-32, merge_ids_op.h:30-32).
 *   split_ids (LoDTensor branch, split_ids_op.h:36-58): out_ids = ids grouped by id % shard_num, input
 *     order and duplicates kept inside a shard; bounds[g] = first position of shard g, bounds[shard_num] =
 *     number of in-range ids (device int64[shard_num+1]).  The workspace keeps the permutation.
 *   merge_ids (merge_ids_op.h:63-87): x = the shards' row blocks laid out shard after shard ([n, d]);
 *     out[i,:] = the next unused row of shard ids[i] % shard_num — needs the workspace of split_ids on the
 *     same ids. */
int b200emb_split_ids(const int64_t* ids, int64_t n, int64_t height, int64_t shard_num, int64_t* out_ids,
                      int64_t* bounds, void* workspace, size_t workspace_bytes, int32_t* status,
                      void* stream);
int b200emb_merge_ids_f32(const float* x, int64_t n, int64_t d, int64_t shard_num, void* workspace,
                          size_t workspace_bytes, float* out, void* stream);
int b200emb_merge_ids_f64(const double* x, int64_t n, int64_t d, int64_t shard_num, void* workspace,
                          size_t workspace_bytes, double* out, void* stream);

/* split_selected_rows — replaces operators/split_selected_rows_op.h:24-102 (whose GPU path issues one
 * cudaMemcpyAsync per row, :83-99): range split of a SelectedRows by `height_sections` (HOST array of
 * n_sections <= 64 section heights, the op's attribute).  Section of a row = FindOutIdx (:24-32): the
 * first section whose successor starts above it; rows past the last start land in the last section.
 * Outputs are laid out section after section, input order kept inside a section:
 *   out_rows[i]  = row - start of its section, out_value[i,:] = its value row,
 *   bounds[s]    = first output position of section s, bounds[n_sections] = n (device int64[n_sections+1]).
 * out_rows doubles as scratch and must not alias rows. */
int b200emb_split_selected_rows_f32(const int64_t* rows, const float* value, int64_t n, int64_t d,
                                    const int64_t* height_sections, int64_t n_sections,
                                    int64_t* out_rows, float* out_value, int64_t* bounds,
                                    void* workspace, size_t workspace_bytes, void* stream);
int b200emb_split_selected_rows_f64(const int64_t* rows, const double* value, int64_t n, int64_t d,
                                    const int64_t* height_sections, int64_t n_sections,
                                    int64_t* out_rows, double* out_value, int64_t* bounds,
                                    void* workspace, size_t workspace_bytes, void* stream);

/* extract_rows — replaces operators/extract_rows_op.cc:49-77: Out[n,1] = the SelectedRows' row ids. */
int b200emb_extract_rows(const int64_t* rows, int64_t n, int64_t* out, void* stream);

/* The sparse weight-decay chain the reference's regularizers append for a SelectedRows gradient
 * (python/paddle/fluid/regularizer.py:146-170 L2Decay, :226-253 L1Decay, added to the gradient at
 * :64-69) — extract_rows -> lookup_table -> [sign] -> scale -> elementwise_add — as ONE kernel:
 *   out[i,:] = addend[i,:] + coeff * f(table[rows[i],:]),  f = identity (mode 0, L2) or sign (mode 1, L1:
 *   Eigen's sign(), operators/sign_op.h:34), every product and sum rounded on its own like the op chain.
 * addend == NULL gives the decay term alone (the chain without the final add); out may alias addend.
 * A row outside [0,height) sets status bit0 (lookup_table_op.h:59-60) and contributes no decay. */
#define B200EMB_DECAY_L2 0
#define B200EMB_DECAY_L1 1
int b200emb_sparse_decay_f32(const float* table, int64_t height, int64_t d, const int64_t* rows,
                             int64_t n, float coeff, int mode, const float* addend, float* out,
                             int32_t* status, void* stream);
int b200emb_sparse_decay_f64(const double* table, int64_t height, int64_t d, const int64_t* rows,
                             int64_t n, double coeff, int mode, const double* addend, double* out,
                             int32_t* status, void* stream);

/* Fused compute + exchange for the row-sharded path (no counterpart in the reference, whose
 * multi-device SelectedRows path copies whole blocks with cudaMemcpyPeer: memory/memcpy.cc:65-72).
 * The destinations are device pointers that may live on OTHER GPUs of the box (peer-mapped, e.g.
 * cudaIpcOpenMemHandle): the kernels store over NVLink while they compute, so no staging buffer and no
 * separate all-to-all are needed; the caller provides the cross-GPU barrier afterwards.
 *   lookup_table_scatter: row k of segment g (seg_start[g] <= k < seg_start[g+1], `segs` segments, device
 *     arrays) of the gathered rows is stored at seg_dst[g] + (k - seg_start[g]) * d  — the owner side of the
 *     lookup: gather from the shard straight into each requester's buffer.
 *   merge_add_scatter: merged row u is stored at shard_dst[g] + (u - shard_bounds[g]) * d for the shard g
 *     with shard_bounds[g] <= u < shard_bounds[g+1]  — the requester side of the update: the local
 *     duplicate reduction writes straight into each owner's receive buffer.  Needs the plan of
 *     b200emb_plan_unique_rows in `workspace` (B200EMB_PLAN_READY) or sorts raw rows otherwise.
 * Destinations must be 16-byte aligned when d*4 is a multiple of 16. */
int b200emb_lookup_table_scatter_f32(const float* table, int64_t height, int64_t d, const int64_t* ids,
                                     int64_t n, int64_t padding_idx, const int64_t* seg_start,
                                     float* const* seg_dst, int64_t segs, int32_t* status, void* stream);
int b200emb_merge_add_scatter_f32(const int64_t* rows, const float* value, int64_t n, int64_t d,
                                  int64_t height, const int64_t* shard_bounds, float* const* shard_dst,
                                  int64_t shards, int64_t* n_unique, void* workspace,
                                  size_t workspace_bytes, unsigned flags, int32_t* status, void* stream);
/*   segment_scatter_i64: dst[g][u - bounds[g]] = src[u] for bounds[g] <= u < bounds[g+1] (bounds[segs] =
 *     element count, read on the device; n_upper = host-side upper bound used to size the grid) — the id
 *     exchange of the lookup: every rank's unique local rows go straight into their owners' buffers. */
int b200emb_segment_scatter_i64(const int64_t* src, int64_t n_upper, const int64_t* bounds,
                                int64_t* const* dst, int64_t segs, void* stream);

/* Set-up utilities for the peer-store kernels above (NOT on the step path; the only entry points that
 * allocate): a device buffer exported through CUDA IPC so that the other ranks' GPUs can store into it,
 * and the import side (cudaIpcOpenMemHandle with lazy peer access).  `handle64` is the 64-byte
 * cudaIpcMemHandle_t, exchanged by the caller (e.g. torch.distributed.all_gather_object). */
int b200emb_peer_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int b200emb_peer_open(const unsigned char* handle64, void** ptr);
int b200emb_peer_close(void* ptr);
int b200emb_peer_free(void* ptr);

/* Stream-ordered barrier across the ranks of the sharded path, over peer memory instead of a
 * collective (no counterpart in the reference).  flag_ptrs: DEVICE array of `world` pointers, entry r
 * = rank r's flag array (uint64[world], zeroed at set-up, peer-mapped like the buffers above).  One
 * tiny kernel: thread r stores `epoch` (release, system scope) into slot [rank] of rank r's array,
 * then waits (acquire) until slot [r] of this rank's own array reaches `epoch`.  `epoch` must grow by
 * one per barrier on every rank.  Everything the preceding kernels of `stream` stored into peer
 * buffers is visible to the kernels the peers launch after their own barrier.  A rank that does not
 * show up within ~20 s raises B200EMB_STATUS_BARRIER_TIMEOUT in `status` instead of hanging. */
int b200emb_peer_barrier(void* const* flag_ptrs, int world, int rank, uint64_t epoch,
                         int32_t* status, void* stream);

/* All-gather of one small int64 row per rank (the split sizes of a step) over peer memory, barrier
 * included — the collective-free counterpart of the all_gather the sharded path needs once per step.
 * table_ptrs: DEVICE array of `world` pointers, entry r = rank r's table (int64[world][cols], peer-mapped);
 * flag_ptrs as in b200emb_peer_barrier but a flag array of its own (uint64[world], zeroed at set-up);
 * local_table = this rank's own table (= entry `rank`).  Thread r stores `row` into slot [rank] of rank r's
 * table, publishes `epoch`, waits for rank r's; then, if host_copy != NULL (pinned host memory,
 * world*cols int64), the complete local table is copied to the host on the same stream. */
int b200emb_peer_allgather_i64(const int64_t* row, int cols, void* const* table_ptrs,
                               const int64_t* local_table, void* const* flag_ptrs, int world, int rank,
                               uint64_t epoch, int64_t* host_copy, int32_t* status, void* stream);

/* Scatter maps of the fused exchange, one tiny kernel instead of a dozen tensor ops on the step path.
 * split_table: device int64 [world][cols], row r = rank r's shard_bounds (b200emb_plan_unique_rows),
 * all-gathered; peer_*_ptrs: device int64[world] = every rank's id / unique-row / gradient receive buffer.
 * With before[g] = sum over ranks r < rank of (split_table[r][g+1] - split_table[r][g]) (this rank's slot in
 * owner g's buffers: ranks before it come first, the reference's gather-concat order) and
 * cnt[r] = split_table[r][rank+1] - split_table[r][rank] (rows rank r requests from this owner):
 *   out[0][g] = peer_i_ptrs[g] + 8*before[g]                    id destinations (segment_scatter_i64)
 *   out[1][r] = cnt[0] + ... + cnt[r-1],  r = 0..world          owner-gather segment starts
 *   out[2][r] = peer_u_ptrs[r] + row_bytes*split_table[r][rank]  owner-gather destinations
 *   out[3][g] = peer_g_ptrs[g] + row_bytes*before[g]             merged-gradient destinations
 * out: device int64 [4][world+1]. */
int b200emb_exchange_maps(const int64_t* split_table, int world, int cols, int rank,
                          const int64_t* peer_i_ptrs, const int64_t* peer_u_ptrs,
                          const int64_t* peer_g_ptrs, int64_t row_bytes, int64_t* out, void* stream);

/* ---- Row-sharded embedding step ------------------------------------------------------------
 * The multi-device SelectedRows path of the reference — ReduceOpHandle::RunImpl gathers every device's sparse
 * gradient on one device (framework/details/reduce_op_handle.cc:88-94, reduce_and_gather.h:56-93), the
 * optimizer runs there and BroadcastOpHandle::RunImpl copies the parameter back
 * (framework/details/broadcast_op_handle.cc:71-122), chosen per gradient by multi_devices_graph_pass.cc:411-426
 * — becomes a table sharded by row over the `world` GPUs of one box: owner(id) = id % world, local row =
 * id / world (the reference's own convention for its distributed lookup table, operators/split_ids_op.h:45-49),
 * rows handed back in the requester's order (operators/merge_ids_op.h:63-87).
 * One engine per rank / GPU; all ranks make the same calls in the same order (SPMD).  The step path never
 * allocates, never synchronises and reads nothing on the host: kernels of different GPUs meet through
 * arrival flags in peer memory, each consumer waiting only for the ranks whose data it reads.  float32, fast
 * (fixed-tree) reduction mode; duplicates are merged locally before they cross NVLink.
 *
 * create   (allocates; current device = this rank's GPU).  `shard` = this rank's rows [ceil(height/world), d]
 *          (row i holds global row i*world + rank), moment1/moment2 alike for Adam (NULL for SGD).
 *          max_ids_per_step bounds n of prepare(); max_recv_rows bounds the unique rows all ranks together
 *          request from one owner in one step.  Both (and world, height, d) must be the SAME on every rank —
 *          the ranks address each other's blocks by layout; connect() returns B200EMB_E_INVALID otherwise.  Exceeding either sets B200EMB_STATUS_SHARD_OVERFLOW in the
 *          status word and drops the excess rows.
 * export / connect   exchange of the per-rank blocks between processes: export() yields a 64-byte CUDA IPC
 *          handle; the caller all-gathers the handles with whatever transport it has (NCCL, MPI,
 *          torch.distributed, Fluid's gen_nccl_id RPC) and passes the world*64 bytes, in rank order, to connect().
 *          Ranks living in ONE process (ParallelExecutor: one thread per device) pass the block pointers
 *          (b200emb_shard_block) to connect_ptrs() instead, with peer access enabled by the caller.
 * prepare  the ids-only part of a step — sort plan, de-duplication, split points per owner, and the hand-over of
 *          the unique row ids to their owners — into one of 4 slots; needs nothing but the ids, so a training
 *          loop can run it for batch i+1 on another stream
 *          while batch i is in flight.  `ids` must stay valid until update() of that slot has run, and
 *          prepare(slot) overwrites the slot: order it (stream wait) after the update() that last used it.
 * lookup   out[k,:] = table[ids[k],:] for the slot's ids (lookup_table_op.h:36-84 over the sharded table).
 * update   sparse SGD / lazy Adam of the sharded table with the SelectedRows gradient {rows = the slot's ids,
 *          value = d_out} (sgd_op.h:47-81, adam_op.h:158-220,282-321): merged locally, merged rows sent to
 *          their owners, fused update there.  lr / beta*_pow are 1-element device tensors.
 * *_phases run a sub-range of the kernel sequence (lookup: 0 nothing — kept for callers of the first layout —,
 *          1 owner gather into the requesters' buffers, 2 fork of the owner-side sort plan, 3 expand; update:
 *          0 merge + send, 1 owner update): a driver that hosts several ranks on ONE GPU issues prepare() and
 *          then phase k for every rank before phase k+1 (kernels that wait for each other must not share a GPU
 *          out of order); everybody else calls lookup()/update().
 * status_ptr  device int32 the kernels OR B200EMB_STATUS_* bits into; read it at your own sync points. */
typedef struct b200emb_shard b200emb_shard_t;
#define B200EMB_SHARD_SGD 0
#define B200EMB_SHARD_ADAM 1
int b200emb_shard_create(int world, int rank, int64_t height, int64_t d, int optimizer, float* shard,
                         float* moment1, float* moment2, int64_t max_ids_per_step, int64_t max_recv_rows,
                         b200emb_shard_t** out);
int b200emb_shard_export(b200emb_shard_t* sh, unsigned char* handle64);
int b200emb_shard_connect(b200emb_shard_t* sh, const unsigned char* handles);
void* b200emb_shard_block(b200emb_shard_t* sh);
int b200emb_shard_connect_ptrs(b200emb_shard_t* sh, void* const* blocks);
int b200emb_shard_prepare(b200emb_shard_t* sh, int slot, const int64_t* ids, int64_t n, void* stream);
int b200emb_shard_lookup(b200emb_shard_t* sh, int slot, float* out, void* stream);
int b200emb_shard_update(b200emb_shard_t* sh, int slot, const float* d_out, const float* lr,
                         const float* beta1_pow, const float* beta2_pow, float beta1, float beta2,
                         float epsilon, void* stream);
int b200emb_shard_lookup_phases(b200emb_shard_t* sh, int slot, float* out, void* stream, int first, int last);
int b200emb_shard_update_phases(b200emb_shard_t* sh, int slot, const float* d_out, const float* lr,
                                const float* beta1_pow, const float* beta2_pow, float beta1, float beta2,
                                float epsilon, void* stream, int first, int last);
int32_t* b200emb_shard_status_ptr(b200emb_shard_t* sh);
/* out8 = {world, rank, shard height, plan height, max ids, per-source id capacity, max recv rows, block bytes} */
int b200emb_shard_info(b200emb_shard_t* sh, int64_t* out8);
/* out6 = device pointers {slot's split points int64[world+1], slot's unique count, owner segment starts
 * int64[world+1], received row ids, step counter u64, status}: for tests and byte accounting after a sync. */
int b200emb_shard_debug_ptrs(b200emb_shard_t* sh, int slot, void** out6);
int b200emb_shard_destroy(b200emb_shard_t* sh);

/* scatter::MergeAdd — replaces math/selected_rows_functor.cu:235-297 (CPU semantics
 * selected_rows_functor.cc:193-228): out_rows = ascending unique rows, out_value[u,:] = sum of
 * the input rows with that id.  out_rows/out_value must hold n rows; *n_unique (device int64)
 * receives U.  */
int b200emb_merge_add_f32(const int64_t* rows, const float* value, int64_t n, int64_t d,
                          int64_t height, int64_t* out_rows, float* out_value,
                          int64_t* n_unique, void* workspace, size_t workspace_bytes,
                          unsigned flags, int32_t* status, void* stream);
int b200emb_merge_add_f64(const int64_t* rows, const double* value, int64_t n, int64_t d,
                          int64_t height, int64_t* out_rows, double* out_value,
                          int64_t* n_unique, void* workspace, size_t workspace_bytes,
                          unsigned flags, int32_t* status, void* stream);

/* SelectedRowsAddToTensor — replaces math/selected_rows_functor.cu:181-226 (CPU
 * selected_rows_functor.cc:153-178): tensor[rows[i],:] += value[i,:], duplicates accumulate. */
int b200emb_selected_rows_add_to_tensor_f32(const int64_t* rows, const float* value, int64_t n,
                                            int64_t d, int64_t height, float* tensor,
                                            void* workspace, size_t workspace_bytes,
                                            unsigned flags, int32_t* status, void* stream);
int b200emb_selected_rows_add_to_tensor_f64(const int64_t* rows, const double* value, int64_t n,
                                            int64_t d, int64_t height, double* tensor,
                                            void* workspace, size_t workspace_bytes,
                                            unsigned flags, int32_t* status, void* stream);

/* sparse SGD (dense Param, SelectedRows Grad), in place — replaces SGDOpCUDAKernel's
 * SelectedRows branch + SparseSGDFunctorKernel (operators/sgd_op.cu:36-53,80-106; CPU
 * semantics sgd_op.h:47-81).  `lr` is the 1-element LearningRate tensor on the device.
 * Duplicate-row reduction is fused with the update: gradient rows are read once and the
 * merged rows never reach memory. */
int b200emb_sparse_sgd_f32(float* param, int64_t height, int64_t d, const int64_t* rows,
                           const float* grad, int64_t n, const float* lr, void* workspace,
                           size_t workspace_bytes, unsigned flags, int32_t* status,
                           void* stream);
int b200emb_sparse_sgd_f64(double* param, int64_t height, int64_t d, const int64_t* rows,
                           const double* grad, int64_t n, const double* lr, void* workspace,
                           size_t workspace_bytes, unsigned flags, int32_t* status,
                           void* stream);

/* sparse (lazy) Adam, in place — replaces AdamOpKernel's SelectedRows branch =
 * scatter::MergeAdd + SparseAdamFunctor via ForRange (operators/adam_op.h:158-220,282-321).
 * lr, beta1_pow, beta2_pow are the 1-element device tensors; beta1/beta2/epsilon are the
 * float attrs (adam_op.cc:99-112), cast to T as adam_op.h:233-235 does. */
int b200emb_sparse_adam_f32(float* param, float* moment1, float* moment2, int64_t height,
                            int64_t d, const int64_t* rows, const float* grad, int64_t n,
                            const float* lr, const float* beta1_pow, const float* beta2_pow,
                            float beta1, float beta2, float epsilon, void* workspace,
                            size_t workspace_bytes, unsigned flags, int32_t* status,
                            void* stream);
int b200emb_sparse_adam_f64(double* param, double* moment1, double* moment2, int64_t height,
                            int64_t d, const int64_t* rows, const double* grad, int64_t n,
                            const double* lr, const double* beta1_pow, const double* beta2_pow,
                            float beta1, float beta2, float epsilon, void* workspace,
                            size_t workspace_bytes, unsigned flags, int32_t* status,
                            void* stream);

/* sparse Adagrad, in place — replaces SparseAdagradFunctor (operators/adagrad_op.cu:26-108;
 * CPU semantics adagrad_op.cc:108-141): merged = MergeAdd(grad); moment[r] += merged^2;
 * param[r] -= lr * merged / (sqrt(moment[r]) + epsilon), fused in one pass like sgd/adam. */
int b200emb_sparse_adagrad_f32(float* param, float* moment, int64_t height, int64_t d,
                               const int64_t* rows, const float* grad, int64_t n, const float* lr,
                               float epsilon, void* workspace, size_t workspace_bytes, unsigned flags,
                               int32_t* status, void* stream);
int b200emb_sparse_adagrad_f64(double* param, double* moment, int64_t height, int64_t d,
                               const int64_t* rows, const double* grad, int64_t n, const double* lr,
                               float epsilon, void* workspace, size_t workspace_bytes, unsigned flags,
                               int32_t* status, void* stream);

/* Sparse SGD without a workspace argument — NO COUNTERPART IN THE REFERENCE (its kernels take
 * temporaries from Fluid's allocator).  Same result as b200emb_sparse_sgd_f32; the scratch is a
 * fixed array of B200EMB_NOALLOC_SCRATCH_BYTES inside the library's device image, so the call
 * never allocates.  Meant for callers that hold nothing but the op's inputs and outputs and must
 * not call cudaMalloc: thin FFI shims, processes whose framework pool already owns the GPU's
 * memory, loops that cannot take cudaMalloc's device-wide sync.
 *   - limit: b200emb_workspace_bytes(n, d, 4) must be <= B200EMB_NOALLOC_SCRATCH_BYTES (8 MiB; e.g.
 *     131072 rows x 64 floats or 65536 x 128), otherwise B200EMB_E_SCRATCH_LIMIT and no launch;
 *   - the array is shared by every call on the device and is not cleaned afterwards: calls that
 *     may overlap on the GPU (other streams / threads) race — serialise them;
 *   - the first call on each device resolves the array's device address
 *     (cudaGetSymbolAddress); make that call before any stream capture;
 *   - the current device must be the one that owns `param`, the other tensors and `stream` (the scratch used
 *     is the current device's copy of the array); a `param` on another device is refused (B200EMB_E_INVALID);
 *   - B200EMB_PLAN_READY is rejected (B200EMB_E_INVALID); B200EMB_DETERMINISTIC is honoured. */
#define B200EMB_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
int b200emb_sparse_sgd_noalloc_f32(float* param, int64_t height, int64_t d, const int64_t* rows,
                                   const float* grad, int64_t n, const float* lr, unsigned flags,
                                   int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200EMB_H_ */
