// capi.cu — the extern "C" boundary declared in include/b200emb.h.
// Argument checks, workspace carving and launch sequencing only; kernels live in
// gather.cu / sort.cu / segreduce.cu.  Nothing here allocates or synchronises.
#include "../../include/b200emb.h"

#include "common.cuh"
#include "internal.h"

using namespace b200emb;

namespace {
thread_local int g_last_cuda_error = 0;

int fail_cuda(cudaError_t e) {
  g_last_cuda_error = static_cast<int>(e);
  return B200EMB_E_CUDA;
}

int check_sizes(int64_t n, int64_t d, int64_t height) {
  if (n < 0 || d <= 0 || height <= 0) return B200EMB_E_INVALID;
  if (n >= (1LL << 30) || height >= 0xFFFFFFFFLL || d >= (1LL << 31)) return B200EMB_E_UNSUPPORTED;
  return B200EMB_OK;
}

int check_workspace(const void* ws, size_t ws_bytes, const Workspace& L) {
  if (ws == nullptr || (reinterpret_cast<uintptr_t>(ws) & 255u) != 0) return B200EMB_E_WORKSPACE;
  if (ws_bytes < L.total) return B200EMB_E_WORKSPACE;
  return B200EMB_OK;
}

// Shared driver of merge_add / add_to_tensor / sgd / adam.
template <typename T>
int run_update(const int64_t* rows, const T* value, int64_t n, int64_t d, int64_t height,
               UpdateArgs<T> args, int64_t* out_rows, int64_t* n_unique, void* workspace,
               size_t workspace_bytes, unsigned flags, int32_t* status, void* stream,
               bool pointers_ok) {
  int rc = check_sizes(n, d, height);
  if (rc != B200EMB_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n == 0) {  // empty SelectedRows is legal everywhere (sgd_op.h:56-58, adam_op.h:285-288)
    if (n_unique) {
      cudaError_t e = cudaMemsetAsync(n_unique, 0, sizeof(int64_t), s);
      if (e != cudaSuccess) return fail_cuda(e);
    }
    return B200EMB_OK;
  }
  if (!pointers_ok || rows == nullptr || value == nullptr || args.param == nullptr)
    return B200EMB_E_INVALID;
  const Workspace L = make_workspace(n, d, sizeof(T));
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  char* ws = static_cast<char*>(workspace);
  const bool det = (flags & B200EMB_DETERMINISTIC) != 0;
  SortedPlan plan;
  cudaError_t e;
  if (flags & B200EMB_PLAN_READY) {
    plan = plan_view(n, height, ws, L);
  } else {
    e = launch_plan(rows, n, height, ws, L, status, s, &plan);
    if (e != cudaSuccess) return fail_cuda(e);
  }
  if (det || args.op == UpdateOp::kMergeAdd) {
    const bool have_unique = (flags & B200EMB_PLAN_READY) && (flags & B200EMB_UNIQUE_READY);
    if (!have_unique) {
      e = launch_head_scan(plan, n, static_cast<uint32_t>(height), ws, L, out_rows, n_unique, s);
      if (e != cudaSuccess) return fail_cuda(e);
    }
    args.uidx = reinterpret_cast<const uint32_t*>(ws + L.off_uidx);
  }
  e = launch_update<T>(plan, value, n, d, height, args, det, ws, L, status, s);
  if (e != cudaSuccess) return fail_cuda(e);
  return B200EMB_OK;
}

template <typename T>
int run_lookup(const T* table, int64_t height, int64_t d, const int64_t* ids, int64_t n,
               int64_t padding_idx, T* out, int32_t* status, void* stream) {
  int rc = check_sizes(n, d, height);
  if (rc != B200EMB_OK) return rc;
  if (n == 0) return B200EMB_OK;
  if (!table || !ids || !out) return B200EMB_E_INVALID;
  cudaError_t e = launch_gather<T>(table, height, d, ids, n, padding_idx, out, status,
                                   static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

template <typename T>
int run_grad_sparse(const int64_t* ids, const T* d_out, int64_t n, int64_t d, int64_t* rows,
                    T* value, void* stream) {
  if (n < 0 || d <= 0) return B200EMB_E_INVALID;
  if (n == 0) return B200EMB_OK;
  if (!ids || !d_out || !rows || !value) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaSuccess;
  if (rows != ids) e = cudaMemcpyAsync(rows, ids, sizeof(int64_t) * n, cudaMemcpyDeviceToDevice, s);
  if (e == cudaSuccess && value != d_out)
    e = cudaMemcpyAsync(value, d_out, sizeof(T) * n * d, cudaMemcpyDeviceToDevice, s);
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}
}  // namespace

extern "C" {

int b200emb_abi_version(void) { return 1; }
int b200emb_last_cuda_error(void) { return g_last_cuda_error; }

size_t b200emb_workspace_bytes(int64_t n, int64_t d, int64_t elem_size) {
  if (n < 0 || d <= 0 || (elem_size != 4 && elem_size != 8)) return 0;
  return make_workspace(n, d, elem_size).total;
}

int b200emb_lookup_table_f32(const float* table, int64_t height, int64_t d, const int64_t* ids,
                             int64_t n, int64_t padding_idx, float* out, int32_t* status,
                             void* stream) {
  return run_lookup<float>(table, height, d, ids, n, padding_idx, out, status, stream);
}
int b200emb_lookup_table_f64(const double* table, int64_t height, int64_t d, const int64_t* ids,
                             int64_t n, int64_t padding_idx, double* out, int32_t* status,
                             void* stream) {
  return run_lookup<double>(table, height, d, ids, n, padding_idx, out, status, stream);
}

int b200emb_lookup_table_grad_sparse_f32(const int64_t* ids, const float* d_out, int64_t n,
                                         int64_t d, int64_t* rows, float* value, void* stream) {
  return run_grad_sparse<float>(ids, d_out, n, d, rows, value, stream);
}
int b200emb_lookup_table_grad_sparse_f64(const int64_t* ids, const double* d_out, int64_t n,
                                         int64_t d, int64_t* rows, double* value, void* stream) {
  return run_grad_sparse<double>(ids, d_out, n, d, rows, value, stream);
}

int b200emb_selected_rows_add_to(const int64_t* in_rows, const void* in_value, int64_t n_in,
                                 int64_t d, int64_t elem_size, int64_t* out_rows,
                                 void* out_value, int64_t row_offset, void* stream) {
  if (n_in < 0 || d <= 0 || row_offset < 0 || elem_size <= 0) return B200EMB_E_INVALID;
  if (n_in == 0) return B200EMB_OK;  // empty inputs are skipped (sum_op.h:129-131)
  if (!in_rows || !in_value || !out_rows || !out_value) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemcpyAsync(out_rows + row_offset, in_rows, sizeof(int64_t) * n_in,
                                  cudaMemcpyDeviceToDevice, s);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(static_cast<char*>(out_value) + row_offset * d * elem_size, in_value,
                        static_cast<size_t>(n_in) * d * elem_size, cudaMemcpyDeviceToDevice, s);
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_plan_rows(const int64_t* rows, int64_t n, int64_t height, void* workspace,
                      size_t workspace_bytes, int32_t* status, void* stream) {
  int rc = check_sizes(n, 1, height);
  if (rc != B200EMB_OK) return rc;
  if (n == 0) return B200EMB_OK;
  if (!rows) return B200EMB_E_INVALID;
  // the plan region does not depend on d / elem_size; only the carry tail does
  const Workspace L = make_workspace(n, 1, 4);
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  cudaError_t e = launch_plan(rows, n, height, static_cast<char*>(workspace), L, status,
                              static_cast<cudaStream_t>(stream), nullptr);
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_plan_unique_rows(const int64_t* rows, int64_t n, int64_t height, int64_t num_shards,
                             int64_t* unique_keys, int64_t* n_unique, int64_t* shard_bounds,
                             int64_t* elem_bounds, int64_t* local_rows, void* workspace,
                             size_t workspace_bytes, int32_t* status, void* stream) {
  int rc = check_sizes(n, 1, height);
  if (rc != B200EMB_OK) return rc;
  if (num_shards < 1 || num_shards > 65536) return B200EMB_E_INVALID;
  const int64_t shard_height = ceil_div(height, num_shards);
  const int64_t plan_height = num_shards > 1 ? num_shards * shard_height : height;
  if (plan_height >= 0xFFFFFFFFLL) return B200EMB_E_UNSUPPORTED;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n > 0 && !rows) return B200EMB_E_INVALID;
  if (n == 0) {
    cudaError_t e = n_unique ? cudaMemsetAsync(n_unique, 0, sizeof(int64_t), s) : cudaSuccess;
    if (e == cudaSuccess && shard_bounds)
      e = cudaMemsetAsync(shard_bounds, 0, sizeof(int64_t) * (num_shards + 1), s);
    if (e == cudaSuccess && elem_bounds)
      e = cudaMemsetAsync(elem_bounds, 0, sizeof(int64_t) * (num_shards + 1), s);
    return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
  }
  const Workspace L = make_workspace(n, 1, 4);
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  char* ws = static_cast<char*>(workspace);
  KeyMap km;
  km.shards = static_cast<uint32_t>(num_shards);
  km.shard_height = static_cast<uint32_t>(shard_height);
  km.raw_height = static_cast<uint32_t>(height);
  SortedPlan plan;
  cudaError_t e = launch_plan(rows, n, plan_height, ws, L, status, s, &plan, &km);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_head_scan(plan, n, static_cast<uint32_t>(plan_height), ws, L, unique_keys, n_unique, s);
  if (e != cudaSuccess) return fail_cuda(e);
  // batch position -> unique index, kept in the position buffer the sort no longer needs (expand_unique)
  e = launch_inverse_map(plan, reinterpret_cast<const uint32_t*>(ws + L.off_uidx), n,
                         reinterpret_cast<uint32_t*>(ws + L.off_perm[(plan.passes & 1) ^ 1]), s);
  if (e != cudaSuccess) return fail_cuda(e);
  if (shard_bounds || elem_bounds || local_rows) {
    if (!unique_keys) return B200EMB_E_INVALID;   // the split reads the unique keys
    e = launch_shard_split(unique_keys, plan, ws, L, static_cast<int>(num_shards),
                           static_cast<uint32_t>(num_shards > 1 ? shard_height : plan_height), n,
                           static_cast<uint32_t>(plan_height), shard_bounds, elem_bounds, local_rows, s);
    if (e != cudaSuccess) return fail_cuda(e);
  }
  return B200EMB_OK;
}

extern "C++" {
namespace {
template <typename T>
int run_permute(const T* src, int64_t n, int64_t d, int64_t plan_height, void* workspace,
                size_t workspace_bytes, T* out, int mode, void* stream) {
  int rc = check_sizes(n, d, plan_height);
  if (rc != B200EMB_OK) return rc;
  if (n == 0) return B200EMB_OK;
  if (!src || !out) return B200EMB_E_INVALID;
  const Workspace L = make_workspace(n, 1, 4);
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  char* ws = static_cast<char*>(workspace);
  SortedPlan plan = plan_view(n, plan_height, ws, L);
  if (mode == 1) {   // expand_unique: out[i,:] = src[inv[i],:], a gather by the inverse map of plan_unique_rows
    plan.perm = reinterpret_cast<const uint32_t*>(ws + L.off_perm[(plan.passes & 1) ^ 1]);
    mode = 0;
  }
  cudaError_t e = launch_permute_rows<T>(src, plan, reinterpret_cast<const uint32_t*>(ws + L.off_uidx), n,
                                         d, out, mode, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}
}  // namespace
}  // extern "C++"

int b200emb_expand_unique_f32(const float* unique_values, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, float* out, void* stream) {
  return run_permute<float>(unique_values, n, d, plan_height, workspace, workspace_bytes, out, 1, stream);
}
int b200emb_expand_unique_f64(const double* unique_values, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, double* out, void* stream) {
  return run_permute<double>(unique_values, n, d, plan_height, workspace, workspace_bytes, out, 1, stream);
}
int b200emb_gather_sorted_f32(const float* value, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, float* out, void* stream) {
  return run_permute<float>(value, n, d, plan_height, workspace, workspace_bytes, out, 0, stream);
}
int b200emb_gather_sorted_f64(const double* value, int64_t n, int64_t d, int64_t plan_height,
                              void* workspace, size_t workspace_bytes, double* out, void* stream) {
  return run_permute<double>(value, n, d, plan_height, workspace, workspace_bytes, out, 0, stream);
}

// ---- split_ids / merge_ids with the reference's exact (duplicate-keeping) semantics ---------------
int b200emb_split_ids(const int64_t* ids, int64_t n, int64_t height, int64_t shard_num, int64_t* out_ids,
                      int64_t* bounds, void* workspace, size_t workspace_bytes, int32_t* status,
                      void* stream) {
  int rc = check_sizes(n, 1, height);
  if (rc != B200EMB_OK) return rc;
  if (shard_num < 1 || shard_num > 65536 || !bounds) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n == 0) {
    cudaError_t e = cudaMemsetAsync(bounds, 0, sizeof(int64_t) * (shard_num + 1), s);
    return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
  }
  if (!ids || !out_ids) return B200EMB_E_INVALID;
  const Workspace L = make_workspace(n, 1, 4);
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  char* ws = static_cast<char*>(workspace);
  KeyMap km;
  km.shards = static_cast<uint32_t>(shard_num);
  km.shard_height = 1;
  km.raw_height = static_cast<uint32_t>(height);
  km.owner_only = true;
  SortedPlan plan;
  cudaError_t e = launch_plan(ids, n, shard_num, ws, L, status, s, &plan, &km);   // stable partition by owner
  if (e != cudaSuccess) return fail_cuda(e);
  // owners present (unique keys) go to out_ids[0..U) for a moment; the gather below overwrites them
  e = launch_head_scan(plan, n, static_cast<uint32_t>(shard_num), ws, L, out_ids, nullptr, s);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_shard_split(out_ids, plan, ws, L, static_cast<int>(shard_num), 1, n,
                         static_cast<uint32_t>(shard_num), nullptr, bounds, nullptr, s);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_permute_rows<double>(reinterpret_cast<const double*>(ids), plan,
                                  reinterpret_cast<const uint32_t*>(ws + L.off_uidx), n, 1,
                                  reinterpret_cast<double*>(out_ids), 0, s);      // 8-byte bit copies
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_merge_ids_f32(const float* x, int64_t n, int64_t d, int64_t shard_num, void* workspace,
                          size_t workspace_bytes, float* out, void* stream) {
  return run_permute<float>(x, n, d, shard_num, workspace, workspace_bytes, out, 2, stream);
}
int b200emb_merge_ids_f64(const double* x, int64_t n, int64_t d, int64_t shard_num, void* workspace,
                          size_t workspace_bytes, double* out, void* stream) {
  return run_permute<double>(x, n, d, shard_num, workspace, workspace_bytes, out, 2, stream);
}

// ---- split_selected_rows / extract_rows / sparse weight decay (regularize.cu) ---------------------
}  // extern "C"
namespace {
template <typename T>
int run_split_selected_rows(const int64_t* rows, const T* value, int64_t n, int64_t d,
                            const int64_t* height_sections, int64_t n_sections, int64_t* out_rows,
                            T* out_value, int64_t* bounds, void* workspace, size_t workspace_bytes,
                            void* stream) {
  if (n_sections < 1 || n_sections > max_sections() || !height_sections || !bounds)
    return B200EMB_E_INVALID;
  int rc = check_sizes(n, d, n_sections);
  if (rc != B200EMB_OK) return rc;
  for (int64_t j = 0; j < n_sections; ++j)
    if (height_sections[j] < 0) return B200EMB_E_INVALID;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n == 0) {
    cudaError_t e = cudaMemsetAsync(bounds, 0, sizeof(int64_t) * (n_sections + 1), s);
    return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
  }
  if (!rows || !value || !out_rows || !out_value || out_rows == rows) return B200EMB_E_INVALID;
  const Workspace L = make_workspace(n, 1, 4);
  rc = check_workspace(workspace, workspace_bytes, L);
  if (rc != B200EMB_OK) return rc;
  char* ws = static_cast<char*>(workspace);
  const int ns = static_cast<int>(n_sections);
  cudaError_t e = launch_section_of(rows, n, height_sections, ns, out_rows, s);   // out_rows = scratch
  if (e != cudaSuccess) return fail_cuda(e);
  SortedPlan plan;
  e = launch_plan(out_rows, n, n_sections, ws, L, nullptr, s, &plan);             // stable partition
  if (e != cudaSuccess) return fail_cuda(e);
  // sections present (unique keys) pass through bounds' neighbour out_rows for a moment; rebase overwrites
  e = launch_head_scan(plan, n, static_cast<uint32_t>(n_sections), ws, L, out_rows, nullptr, s);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_shard_split(out_rows, plan, ws, L, ns, 1, n, static_cast<uint32_t>(n_sections), nullptr,
                         bounds, nullptr, s);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_rebase_rows(rows, plan, n, height_sections, ns, out_rows, s);
  if (e != cudaSuccess) return fail_cuda(e);
  e = launch_permute_rows<T>(value, plan, reinterpret_cast<const uint32_t*>(ws + L.off_uidx), n, d,
                             out_value, 0, s);
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

template <typename T>
int run_sparse_decay(const T* table, int64_t height, int64_t d, const int64_t* rows, int64_t n, T coeff,
                     int mode, const T* addend, T* out, int32_t* status, void* stream) {
  int rc = check_sizes(n, d, height);
  if (rc != B200EMB_OK) return rc;
  if (mode != B200EMB_DECAY_L2 && mode != B200EMB_DECAY_L1) return B200EMB_E_INVALID;
  if (n == 0) return B200EMB_OK;
  if (!table || !rows || !out) return B200EMB_E_INVALID;
  cudaError_t e = launch_sparse_decay<T>(table, height, d, rows, n, coeff, mode, addend, out, status,
                                         static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}
}  // namespace
extern "C" {

int b200emb_split_selected_rows_f32(const int64_t* rows, const float* value, int64_t n, int64_t d,
                                    const int64_t* height_sections, int64_t n_sections,
                                    int64_t* out_rows, float* out_value, int64_t* bounds,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  return run_split_selected_rows<float>(rows, value, n, d, height_sections, n_sections, out_rows,
                                        out_value, bounds, workspace, workspace_bytes, stream);
}
int b200emb_split_selected_rows_f64(const int64_t* rows, const double* value, int64_t n, int64_t d,
                                    const int64_t* height_sections, int64_t n_sections,
                                    int64_t* out_rows, double* out_value, int64_t* bounds,
                                    void* workspace, size_t workspace_bytes, void* stream) {
  return run_split_selected_rows<double>(rows, value, n, d, height_sections, n_sections, out_rows,
                                         out_value, bounds, workspace, workspace_bytes, stream);
}

int b200emb_extract_rows(const int64_t* rows, int64_t n, int64_t* out, void* stream) {
  if (n < 0) return B200EMB_E_INVALID;
  if (n == 0) return B200EMB_OK;
  if (!rows || !out) return B200EMB_E_INVALID;
  if (rows == out) return B200EMB_OK;
  cudaError_t e = cudaMemcpyAsync(out, rows, sizeof(int64_t) * n, cudaMemcpyDeviceToDevice,
                                  static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_sparse_decay_f32(const float* table, int64_t height, int64_t d, const int64_t* rows,
                             int64_t n, float coeff, int mode, const float* addend, float* out,
                             int32_t* status, void* stream) {
  return run_sparse_decay<float>(table, height, d, rows, n, coeff, mode, addend, out, status, stream);
}
int b200emb_sparse_decay_f64(const double* table, int64_t height, int64_t d, const int64_t* rows,
                             int64_t n, double coeff, int mode, const double* addend, double* out,
                             int32_t* status, void* stream) {
  return run_sparse_decay<double>(table, height, d, rows, n, coeff, mode, addend, out, status, stream);
}

// ---- peer-store variants for the sharded path (destinations may be other GPUs' memory) ----------
int b200emb_lookup_table_scatter_f32(const float* table, int64_t height, int64_t d, const int64_t* ids,
                                     int64_t n, int64_t padding_idx, const int64_t* seg_start,
                                     float* const* seg_dst, int64_t segs, int32_t* status, void* stream) {
  int rc = check_sizes(n, d, height);
  if (rc != B200EMB_OK) return rc;
  if (n == 0) return B200EMB_OK;
  if (!table || !ids || !seg_start || !seg_dst || segs < 1 || segs > 4096) return B200EMB_E_INVALID;
  cudaError_t e = launch_gather<float>(table, height, d, ids, n, padding_idx, nullptr, status,
                                       static_cast<cudaStream_t>(stream), seg_start, seg_dst,
                                       static_cast<int>(segs));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_segment_scatter_i64(const int64_t* src, int64_t n_upper, const int64_t* bounds,
                                int64_t* const* dst, int64_t segs, void* stream) {
  if (n_upper < 0 || segs < 1 || segs > 4096) return B200EMB_E_INVALID;
  if (n_upper == 0) return B200EMB_OK;
  if (!src || !bounds || !dst) return B200EMB_E_INVALID;
  cudaError_t e = launch_segment_scatter_i64(src, n_upper, bounds, dst, static_cast<int>(segs),
                                             static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200EMB_OK : fail_cuda(e);
}

int b200emb_merge_add_scatter_f32(const int64_t* rows, const float* value, int64_t n, int64_t d,
                                  int64_t height, const int64_t* shard_bounds, float* const* shard_dst,
                                  int64_t shards, int64_t* n_unique, void* workspace,
                                  size_t workspace_bytes, unsigned flags, int32_t* status, void* stream) {
  if (shards < 1 || shards > 4096 || !shard_bounds || !shard_dst) return B200EMB_E_INVALID;
  UpdateArgs<float> a;
  a.op = UpdateOp::kMergeAdd;
  a.param = const_cast<float*>(value);   // unused (every row goes through the scatter map); non-null for the checks
  a.shard_bounds = shard_bounds;
  a.shard_dst = shard_dst;
  a.shards = static_cast<int>(shards);
  a.dst_aligned16 = (d % 4) == 0;        // destinations are required to be 16-byte aligned buffers
  return run_update<float>(rows, value, n, d, height, a, /*out_rows=*/nullptr, n_unique, workspace,
                           workspace_bytes, flags, status, stream, true);
}

#define B200EMB_DEFINE_UPDATES(T, S)                                                             \
  int b200emb_merge_add_##S(const int64_t* rows, const T* value, int64_t n, int64_t d,           \
                            int64_t height, int64_t* out_rows, T* out_value, int64_t* n_unique,  \
                            void* workspace, size_t workspace_bytes, unsigned flags,             \
                            int32_t* status, void* stream) {                                     \
    UpdateArgs<T> a;                                                                             \
    a.op = UpdateOp::kMergeAdd;                                                                  \
    a.param = out_value;                                                                         \
    return run_update<T>(rows, value, n, d, height, a, out_rows, n_unique, workspace,            \
                         workspace_bytes, flags, status, stream, out_rows && n_unique);          \
  }                                                                                              \
  int b200emb_selected_rows_add_to_tensor_##S(const int64_t* rows, const T* value, int64_t n,    \
                                              int64_t d, int64_t height, T* tensor,              \
                                              void* workspace, size_t workspace_bytes,           \
                                              unsigned flags, int32_t* status, void* stream) {   \
    UpdateArgs<T> a;                                                                             \
    a.op = UpdateOp::kAddToTensor;                                                               \
    a.param = tensor;                                                                            \
    return run_update<T>(rows, value, n, d, height, a, nullptr, nullptr, workspace,              \
                         workspace_bytes, flags, status, stream, true);                          \
  }                                                                                              \
  int b200emb_sparse_sgd_##S(T* param, int64_t height, int64_t d, const int64_t* rows,           \
                             const T* grad, int64_t n, const T* lr, void* workspace,             \
                             size_t workspace_bytes, unsigned flags, int32_t* status,            \
                             void* stream) {                                                     \
    UpdateArgs<T> a;                                                                             \
    a.op = UpdateOp::kSgd;                                                                       \
    a.param = param;                                                                             \
    a.lr = lr;                                                                                   \
    return run_update<T>(rows, grad, n, d, height, a, nullptr, nullptr, workspace,               \
                         workspace_bytes, flags, status, stream, lr != nullptr);                 \
  }                                                                                              \
  int b200emb_sparse_adam_##S(T* param, T* moment1, T* moment2, int64_t height, int64_t d,       \
                              const int64_t* rows, const T* grad, int64_t n, const T* lr,        \
                              const T* beta1_pow, const T* beta2_pow, float beta1, float beta2,  \
                              float epsilon, void* workspace, size_t workspace_bytes,            \
                              unsigned flags, int32_t* status, void* stream) {                   \
    UpdateArgs<T> a;                                                                             \
    a.op = UpdateOp::kAdam;                                                                      \
    a.param = param;                                                                             \
    a.moment1 = moment1;                                                                         \
    a.moment2 = moment2;                                                                         \
    a.lr = lr;                                                                                   \
    a.beta1_pow = beta1_pow;                                                                     \
    a.beta2_pow = beta2_pow;                                                                     \
    a.beta1 = static_cast<T>(beta1);                                                             \
    a.beta2 = static_cast<T>(beta2);                                                             \
    a.epsilon = static_cast<T>(epsilon);                                                         \
    return run_update<T>(rows, grad, n, d, height, a, nullptr, nullptr, workspace,               \
                         workspace_bytes, flags, status, stream,                                 \
                         lr && beta1_pow && beta2_pow && moment1 && moment2);                    \
  }

#define B200EMB_DEFINE_ADAGRAD(T, S)                                                             \
  int b200emb_sparse_adagrad_##S(T* param, T* moment, int64_t height, int64_t d,                 \
                                 const int64_t* rows, const T* grad, int64_t n, const T* lr,     \
                                 float epsilon, void* workspace, size_t workspace_bytes,         \
                                 unsigned flags, int32_t* status, void* stream) {                \
    UpdateArgs<T> a;                                                                             \
    a.op = UpdateOp::kAdagrad;                                                                   \
    a.param = param;                                                                             \
    a.moment1 = moment;                                                                          \
    a.lr = lr;                                                                                   \
    a.epsilon = static_cast<T>(epsilon);                                                         \
    return run_update<T>(rows, grad, n, d, height, a, nullptr, nullptr, workspace,               \
                         workspace_bytes, flags, status, stream, lr && moment);                  \
  }
B200EMB_DEFINE_ADAGRAD(float, f32)
B200EMB_DEFINE_ADAGRAD(double, f64)

B200EMB_DEFINE_UPDATES(float, f32)
B200EMB_DEFINE_UPDATES(double, f64)

}  // extern "C"
