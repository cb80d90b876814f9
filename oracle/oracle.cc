// oracle.cc — CPU restatement of the reference's sparse-embedding hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product path (paddle-1_b200/, csrc/)
// links, imports or executes this file; it is the checker used by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
//
// Every function restates one reference CPU routine (paths relative to
// /root/reference/paddle/fluid/operators) with the same loop order and the same
// sequence of floating-point operations.  Build with -O2 -ffp-contract=off
// (the reference CPU build is -O3 -mavx, no FMA: cmake/simd.cmake:11), so no
// multiply-add is ever contracted.
//
// Parity is PINNED: (1) by the hand-written constants of the reference's own
// tests, restated as tests/golden/*.json (G1..G10 of SURVEY.md §8c, G11), and
// (2) in the build container by oracle/_ref, the reference's unmodified CPU
// kernels compiled from /root/reference (see oracle/Makefile, tests/test_ref_*).
//
// Adam interpretation: `sqrt` is the float overload,
// every sub-expression stays in T, `(1-beta2)*g*g` associates left to right.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <numeric>
#include <set>
#include <unordered_map>
#include <vector>

namespace {

// ---------------------------------------------------------------------------
// lookup_table forward: lookup_table_op.h:36-84 (LoDTensor table branch 46-63)
// returns 0, or -(k+1) when ids[k] violates the reference's PADDLE_ENFORCE range.
template <typename T>
int64_t LookupTable(const T* table, int64_t height, int64_t width,
                    const int64_t* ids, int64_t n, int64_t padding_idx, T* out) {
  constexpr int64_t kNoPadding = -1;  // lookup_table_op.h:33
  for (int64_t k = 0; k < n; ++k) {
    T* dst = out + k * width;
    if (padding_idx != kNoPadding && ids[k] == padding_idx) {
      std::memset(dst, 0, sizeof(T) * width);           // :56-57
    } else {
      if (!(ids[k] < height) || !(ids[k] >= 0)) return -(k + 1);  // :59-60
      std::memcpy(dst, table + ids[k] * width, sizeof(T) * width);  // :61-62
    }
  }
  return 0;
}

// lookup_table with a SelectedRows table: lookup_table_op.h:64-81; Index() is a
// linear first-match search (framework/selected_rows.h:83-89).
template <typename T>
int64_t LookupTableSelectedRows(const int64_t* table_rows, int64_t n_table_rows,
                                const T* table_value, int64_t width,
                                const int64_t* ids, int64_t n,
                                int64_t padding_idx, T* out) {
  constexpr int64_t kNoPadding = -1;
  for (int64_t k = 0; k < n; ++k) {
    T* dst = out + k * width;
    if (padding_idx != kNoPadding && ids[k] == padding_idx) {
      std::memset(dst, 0, sizeof(T) * width);
    } else {
      if (ids[k] < 0) return -(k + 1);
      const int64_t* hit = std::find(table_rows, table_rows + n_table_rows, ids[k]);
      if (hit == table_rows + n_table_rows) return -(k + 1);
      std::memcpy(dst, table_value + (hit - table_rows) * width, sizeof(T) * width);
    }
  }
  return 0;
}

// lookup_table_grad, sparse branch: lookup_table_op.h:106-134.
// rows = ids verbatim (duplicates kept, input order); value = copy of dOut.
template <typename T>
void LookupTableGradSparse(const int64_t* ids, int64_t n, const T* d_out,
                           int64_t width, int64_t* rows, T* value) {
  for (int64_t i = 0; i < n; ++i) rows[i] = ids[i];        // :114-118
  std::memcpy(value, d_out, sizeof(T) * n * width);          // :133
}

// lookup_table_grad, dense branch: lookup_table_op.h:135-157.
template <typename T>
int64_t LookupTableGradDense(const int64_t* ids, int64_t n, const T* d_out,
                             int64_t height, int64_t width, T* d_table) {
  std::memset(d_table, 0, sizeof(T) * height * width);      // :147
  for (int64_t i = 0; i < n; ++i) {
    if (!(ids[i] < height) || !(ids[i] >= 0)) return -(i + 1);
    for (int64_t j = 0; j < width; ++j)
      d_table[ids[i] * width + j] += d_out[i * width + j];  // :152-154
  }
  return 0;
}

// SelectedRowsAddTo: math/selected_rows_functor.cc:116-146. Appends rows, copies
// the value block at element offset `offset` of the destination value.
template <typename T>
void SelectedRowsAddTo(const int64_t* in_rows, int64_t n_in, const T* in_value,
                       int64_t width, int64_t* out_rows, int64_t out_rows_fill,
                       T* out_value, int64_t offset) {
  for (int64_t i = 0; i < n_in; ++i) out_rows[out_rows_fill + i] = in_rows[i];
  std::memcpy(out_value + offset, in_value, sizeof(T) * n_in * width);
}

// SelectedRowsAddToTensor: math/selected_rows_functor.cc:153-178.
template <typename T>
void SelectedRowsAddToTensor(const int64_t* rows, int64_t n, const T* value,
                             int64_t width, T* tensor) {
  for (int64_t i = 0; i < n; ++i)
    for (int64_t j = 0; j < width; ++j)
      tensor[rows[i] * width + j] += value[i * width + j];
}

// SelectedRowsAddTensor: math/selected_rows_functor.cc:77-114
// out = 0; out[rows[i]] += value[i]; out = out + in2.
template <typename T>
void SelectedRowsAddTensor(const int64_t* rows, int64_t n, const T* value,
                           int64_t width, int64_t height, const T* in2, T* out) {
  for (int64_t e = 0; e < height * width; ++e) out[e] = static_cast<T>(0.0);
  SelectedRowsAddToTensor<T>(rows, n, value, width, out);
  for (int64_t e = 0; e < height * width; ++e) out[e] = out[e] + in2[e];
}

// scatter::MergeAdd, reference form: math/selected_rows_functor.cc:193-228.
// out_rows = ascending unique (std::set); out zero-filled; then for every input
// row IN INPUT ORDER a linear FindPos (:193-195) and a column loop of `+=`.
// Returns U.
template <typename T>
int64_t MergeAddRef(const int64_t* rows, int64_t n, const T* value, int64_t width,
                    int64_t* out_rows, T* out_value) {
  std::set<int64_t> row_set(rows, rows + n);
  std::vector<int64_t> merge_rows(row_set.begin(), row_set.end());
  const int64_t u = static_cast<int64_t>(merge_rows.size());
  for (int64_t i = 0; i < u; ++i) out_rows[i] = merge_rows[i];
  for (int64_t e = 0; e < u * width; ++e) out_value[e] = static_cast<T>(0.0);
  for (int64_t i = 0; i < n; ++i) {
    size_t pos = std::find(merge_rows.begin(), merge_rows.end(), rows[i]) -
                 merge_rows.begin();
    for (int64_t j = 0; j < width; ++j)
      out_value[pos * width + j] += value[i * width + j];
  }
  return u;
}

// Same result bit for bit (same per-row add order: input order, from +0), but the
// position lookup is a hash map so it finishes at BASELINE sizes.
template <typename T>
int64_t MergeAddFast(const int64_t* rows, int64_t n, const T* value, int64_t width,
                     int64_t* out_rows, T* out_value) {
  std::vector<int64_t> merge_rows(rows, rows + n);
  std::sort(merge_rows.begin(), merge_rows.end());
  merge_rows.erase(std::unique(merge_rows.begin(), merge_rows.end()), merge_rows.end());
  const int64_t u = static_cast<int64_t>(merge_rows.size());
  std::unordered_map<int64_t, int64_t> pos_of;
  pos_of.reserve(static_cast<size_t>(u) * 2);
  for (int64_t i = 0; i < u; ++i) { out_rows[i] = merge_rows[i]; pos_of[merge_rows[i]] = i; }
  for (int64_t e = 0; e < u * width; ++e) out_value[e] = static_cast<T>(0.0);
  for (int64_t i = 0; i < n; ++i) {
    const int64_t pos = pos_of[rows[i]];
    for (int64_t j = 0; j < width; ++j)
      out_value[pos * width + j] += value[i * width + j];
  }
  return u;
}

// fp64 "truth" for error accounting: accumulate in double, input order.
template <typename T>
int64_t MergeAddTruth(const int64_t* rows, int64_t n, const T* value, int64_t width,
                      int64_t* out_rows, double* out_value) {
  std::vector<int64_t> merge_rows(rows, rows + n);
  std::sort(merge_rows.begin(), merge_rows.end());
  merge_rows.erase(std::unique(merge_rows.begin(), merge_rows.end()), merge_rows.end());
  const int64_t u = static_cast<int64_t>(merge_rows.size());
  std::unordered_map<int64_t, int64_t> pos_of;
  pos_of.reserve(static_cast<size_t>(u) * 2);
  for (int64_t i = 0; i < u; ++i) { out_rows[i] = merge_rows[i]; pos_of[merge_rows[i]] = i; }
  for (int64_t e = 0; e < u * width; ++e) out_value[e] = 0.0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t pos = pos_of[rows[i]];
    for (int64_t j = 0; j < width; ++j)
      out_value[pos * width + j] += static_cast<double>(value[i * width + j]);
  }
  return u;
}

// sparse SGD (dense param, SelectedRows grad): sgd_op.h:47-81.
// NO merge: one `p -= lr*g` per occurrence, input order, product rounded first.
template <typename T>
int64_t SgdSparse(T* param, int64_t height, int64_t width, const int64_t* rows,
                  int64_t n, const T* grad, const T* lr) {
  if (n == 0) return 0;                                       // :56-58
  for (int64_t i = 0; i < n; ++i) {
    if (!(rows[i] < height)) return -(i + 1);                 // :75-76
    for (int64_t j = 0; j < width; ++j)
      param[rows[i] * width + j] -= lr[0] * grad[i * width + j];  // :77-79
  }
  return 0;
}

// SparseAdamFunctor: adam_op.h:199-219, applied to MergeAdd's output by
// AdamOpKernel's SelectedRows branch (:282-321) through ForRange<CPU>
// (platform/for_range.h:29-41).
template <typename T>
void AdamRows(T* param, T* mom1, T* mom2, int64_t width, const int64_t* mrows,
              int64_t u, const T* mgrad, const T* lr_p, const T* beta1_pow_p,
              const T* beta2_pow_p, T beta1, T beta2, T epsilon) {
  for (int64_t i = 0; i < u; ++i) {
    T beta1_pow = *beta1_pow_p;
    T beta2_pow = *beta2_pow_p;
    for (int64_t j = 0; j < width; ++j) {
      T g = mgrad[i * width + j];
      T m1 = mom1[mrows[i] * width + j];
      T m2 = mom2[mrows[i] * width + j];
      T lr = *lr_p;
      T p = param[mrows[i] * width + j];
      lr *= std::sqrt(1 - beta2_pow) / (1 - beta1_pow);          // :209
      m1 = beta1 * m1 + (1 - beta1) * g;                          // :211
      m2 = beta2 * m2 + (1 - beta2) * g * g;                      // :212
      p -= lr * (m1 / (std::sqrt(m2) + epsilon));                 // :213
      mom1[mrows[i] * width + j] = m1;
      mom2[mrows[i] * width + j] = m2;
      param[mrows[i] * width + j] = p;
    }
  }
}

template <typename T>
int64_t AdamSparse(T* param, T* mom1, T* mom2, int64_t height, int64_t width,
                   const int64_t* rows, int64_t n, const T* grad, const T* lr,
                   const T* beta1_pow, const T* beta2_pow, T beta1, T beta2,
                   T epsilon, int reference_form) {
  (void)height;
  if (n == 0) return 0;                                       // :285-288
  std::vector<int64_t> mrows(n);
  std::vector<T> mgrad(static_cast<size_t>(n) * width);
  int64_t u = reference_form
                  ? MergeAddRef<T>(rows, n, grad, width, mrows.data(), mgrad.data())
                  : MergeAddFast<T>(rows, n, grad, width, mrows.data(), mgrad.data());
  AdamRows<T>(param, mom1, mom2, width, mrows.data(), u, mgrad.data(), lr,
              beta1_pow, beta2_pow, beta1, beta2, epsilon);
  return u;
}

// SparseAdagradFunctor<CPU>: adagrad_op.cc:108-141 (next-row §8f-1).
// merged = MergeAdd(grad); moment[rows] += merged*merged; then
// param[r] -= lr*merged/(sqrt(moment[r])+eps), evaluated left to right.
template <typename T>
int64_t AdagradSparse(T* param, T* moment, int64_t height, int64_t width,
                      const int64_t* rows, int64_t n, const T* grad, const T* lr,
                      T epsilon) {
  (void)height;
  if (n == 0) return 0;
  std::vector<int64_t> mrows(n);
  std::vector<T> mgrad(static_cast<size_t>(n) * width);
  int64_t u = MergeAddFast<T>(rows, n, grad, width, mrows.data(), mgrad.data());
  std::vector<T> sq(static_cast<size_t>(u) * width);
  for (int64_t e = 0; e < u * width; ++e) sq[e] = mgrad[e] * mgrad[e];     // scatter::Mul
  SelectedRowsAddToTensor<T>(mrows.data(), u, sq.data(), width, moment);   // :124-125
  for (int64_t i = 0; i < u; ++i)
    for (int64_t j = 0; j < width; ++j)
      param[mrows[i] * width + j] -=
          lr[0] * mgrad[i * width + j] /
          (std::sqrt(moment[mrows[i] * width + j]) + epsilon);             // :132-138
  return u;
}

}  // namespace

// ---------------------------------------------------------------------------
// C ABI (ctypes).  _f32 / _f64 pairs.
#define ORACLE_FOR_TYPES(X) X(float, f32) X(double, f64)

extern "C" {

#define DEFINE_LOOKUP(T, S)                                                              \
  int64_t oracle_lookup_table_##S(const T* w, int64_t v, int64_t d, const int64_t* ids,  \
                                  int64_t n, int64_t padding_idx, T* out) {              \
    return LookupTable<T>(w, v, d, ids, n, padding_idx, out);                            \
  }                                                                                      \
  int64_t oracle_lookup_table_selected_rows_##S(                                         \
      const int64_t* trows, int64_t ntrows, const T* tvalue, int64_t d,                  \
      const int64_t* ids, int64_t n, int64_t padding_idx, T* out) {                      \
    return LookupTableSelectedRows<T>(trows, ntrows, tvalue, d, ids, n, padding_idx, out); \
  }                                                                                      \
  void oracle_lookup_table_grad_sparse_##S(const int64_t* ids, int64_t n, const T* dout, \
                                           int64_t d, int64_t* rows, T* value) {         \
    LookupTableGradSparse<T>(ids, n, dout, d, rows, value);                              \
  }                                                                                      \
  int64_t oracle_lookup_table_grad_dense_##S(const int64_t* ids, int64_t n,              \
                                             const T* dout, int64_t v, int64_t d,        \
                                             T* dtable) {                                \
    return LookupTableGradDense<T>(ids, n, dout, v, d, dtable);                          \
  }                                                                                      \
  void oracle_selected_rows_add_to_##S(const int64_t* in_rows, int64_t n_in,             \
                                       const T* in_value, int64_t d, int64_t* out_rows,  \
                                       int64_t out_rows_fill, T* out_value,              \
                                       int64_t offset) {                                 \
    SelectedRowsAddTo<T>(in_rows, n_in, in_value, d, out_rows, out_rows_fill, out_value, \
                         offset);                                                        \
  }                                                                                      \
  void oracle_selected_rows_add_to_tensor_##S(const int64_t* rows, int64_t n,            \
                                              const T* value, int64_t d, T* tensor) {    \
    SelectedRowsAddToTensor<T>(rows, n, value, d, tensor);                               \
  }                                                                                      \
  void oracle_selected_rows_add_tensor_##S(const int64_t* rows, int64_t n,               \
                                           const T* value, int64_t d, int64_t height,    \
                                           const T* in2, T* out) {                       \
    SelectedRowsAddTensor<T>(rows, n, value, d, height, in2, out);                       \
  }                                                                                      \
  int64_t oracle_merge_add_ref_##S(const int64_t* rows, int64_t n, const T* value,       \
                                   int64_t d, int64_t* out_rows, T* out_value) {         \
    return MergeAddRef<T>(rows, n, value, d, out_rows, out_value);                       \
  }                                                                                      \
  int64_t oracle_merge_add_##S(const int64_t* rows, int64_t n, const T* value,           \
                               int64_t d, int64_t* out_rows, T* out_value) {             \
    return MergeAddFast<T>(rows, n, value, d, out_rows, out_value);                      \
  }                                                                                      \
  int64_t oracle_merge_add_truth_##S(const int64_t* rows, int64_t n, const T* value,     \
                                     int64_t d, int64_t* out_rows, double* out_value) {  \
    return MergeAddTruth<T>(rows, n, value, d, out_rows, out_value);                     \
  }                                                                                      \
  int64_t oracle_sgd_sparse_##S(T* param, int64_t v, int64_t d, const int64_t* rows,     \
                                int64_t n, const T* grad, const T* lr) {                 \
    return SgdSparse<T>(param, v, d, rows, n, grad, lr);                                 \
  }                                                                                      \
  int64_t oracle_adam_sparse_##S(T* param, T* m1, T* m2, int64_t v, int64_t d,           \
                                 const int64_t* rows, int64_t n, const T* grad,          \
                                 const T* lr, const T* b1pow, const T* b2pow, float b1,  \
                                 float b2, float eps, int reference_form) {              \
    /* attrs are float and cast to T: adam_op.h:233-235 */                               \
    return AdamSparse<T>(param, m1, m2, v, d, rows, n, grad, lr, b1pow, b2pow,           \
                         static_cast<T>(b1), static_cast<T>(b2), static_cast<T>(eps),    \
                         reference_form);                                                \
  }                                                                                      \
  int64_t oracle_adagrad_sparse_##S(T* param, T* moment, int64_t v, int64_t d,           \
                                    const int64_t* rows, int64_t n, const T* grad,       \
                                    const T* lr, float eps) {                            \
    return AdagradSparse<T>(param, moment, v, d, rows, n, grad, lr,                      \
                            static_cast<T>(eps));                                        \
  }
ORACLE_FOR_TYPES(DEFINE_LOOKUP)

// split_ids on a LoDTensor of ids: split_ids_op.h:36-58. shard = id % shard_num,
// order within a shard = input order.  `out` is one buffer of n ids laid out shard
// after shard; counts[s] = ids in shard s.
void oracle_split_ids(const int64_t* ids, int64_t n, int64_t shard_num, int64_t* out,
                      int64_t* counts) {
  std::vector<std::vector<int64_t>> buckets(static_cast<size_t>(shard_num));
  for (int64_t i = 0; i < n; ++i) {
    size_t shard = static_cast<size_t>(ids[i]) % static_cast<size_t>(shard_num);
    buckets[shard].push_back(ids[i]);
  }
  int64_t fill = 0;
  for (int64_t s = 0; s < shard_num; ++s) {
    counts[s] = static_cast<int64_t>(buckets[s].size());
    for (int64_t id : buckets[s]) out[fill++] = id;
  }
}

// merge_ids: merge_ids_op.h:63-87.  x = the shards' row blocks laid out shard after
// shard (x_offsets[s] = first row of shard s); row i of out is the next unused row
// of shard ids[i] % shard_num.
void oracle_merge_ids_f32(const int64_t* ids, int64_t n, int64_t shard_num,
                          const float* x, const int64_t* x_offsets, int64_t d,
                          float* out) {
  std::vector<int64_t> cursor(static_cast<size_t>(shard_num), 0);
  for (int64_t i = 0; i < n; ++i) {
    size_t shard = static_cast<size_t>(ids[i]) % static_cast<size_t>(shard_num);
    int64_t src_row = x_offsets[shard] + cursor[shard];
    std::memcpy(out + i * d, x + src_row * d, sizeof(float) * d);
    cursor[shard] += 1;
  }
}

// split_ids on a SelectedRows: split_ids_op.h:59-92.  Rows go to shard row%S in input
// order; the value row copied for an output row is the LAST input row holding that id
// (id_to_index overwrite, :77).  out_rows / out_value laid out shard after shard.
void oracle_split_ids_selected_rows_f32(const int64_t* rows, int64_t n, const float* value,
                                        int64_t d, int64_t shard_num, int64_t* out_rows,
                                        float* out_value, int64_t* counts) {
  std::unordered_map<int64_t, int64_t> last_index;
  std::vector<std::vector<int64_t>> buckets(static_cast<size_t>(shard_num));
  for (int64_t i = 0; i < n; ++i) {
    last_index[rows[i]] = i;
    buckets[static_cast<size_t>(rows[i]) % static_cast<size_t>(shard_num)].push_back(rows[i]);
  }
  int64_t fill = 0;
  for (int64_t s = 0; s < shard_num; ++s) {
    counts[s] = static_cast<int64_t>(buckets[s].size());
    for (int64_t id : buckets[s]) {
      out_rows[fill] = id;
      std::memcpy(out_value + fill * d, value + last_index[id] * d, sizeof(float) * d);
      ++fill;
    }
  }
}

// split_selected_rows: split_selected_rows_op.h:24-102. Range split by height_sections;
// output rows are rebased to the section start. Laid out section after section.
void oracle_split_selected_rows_f32(const int64_t* rows, int64_t n, const float* value,
                                    int64_t d, const int64_t* height_sections,
                                    int64_t n_sections, int64_t* out_rows,
                                    float* out_value, int64_t* counts) {
  std::vector<int64_t> abs(static_cast<size_t>(n_sections), 0);
  for (int64_t s = 1; s < n_sections; ++s) abs[s] = abs[s - 1] + height_sections[s - 1];
  std::vector<std::vector<int64_t>> idx(static_cast<size_t>(n_sections));
  for (int64_t i = 0; i < n; ++i) {
    int64_t sec = n_sections - 1;
    for (int64_t s = 1; s < n_sections; ++s)
      if (rows[i] < abs[s]) { sec = s - 1; break; }
    idx[sec].push_back(i);
  }
  int64_t fill = 0;
  for (int64_t s = 0; s < n_sections; ++s) {
    counts[s] = static_cast<int64_t>(idx[s].size());
    for (int64_t i : idx[s]) {
      out_rows[fill] = rows[i] - abs[s];
      std::memcpy(out_value + fill * d, value + i * d, sizeof(float) * d);
      ++fill;
    }
  }
}

// Sparse weight decay: the op chain the regularizers append for a SelectedRows gradient
// (python/paddle/fluid/regularizer.py:146-170 L2Decay, :226-253 L1Decay, then :64-69), op by op:
//   extract_rows  (extract_rows_op.cc:49-77)   idx = the gradient's row ids, duplicates kept
//   lookup_table  (lookup_table_op.h:46-63)    decay[i,:] = W[idx[i],:]
//   sign          (sign_op.h:34, L1 only)      Eigen's scalar_sign_op (third-party Eigen, pinned in
//                                              cmake/external/eigen.cmake:27 to 917060c3; its real-type
//                                              functor is Scalar((a > 0) - (a < 0)))
//   scale         (scale_op.h:53-56)           decay = scale * decay
//   elementwise_add(grad, decay)               out = grad + decay       (skipped when addend == null)
// returns 0 or -(k+1) when rows[k] violates lookup_table's range check.
}  // extern "C"
namespace {
template <typename T>
int64_t SparseDecay(const T* table, int64_t height, int64_t width, const int64_t* rows, int64_t n,
                    T coeff, int mode, const T* addend, T* out) {
  std::vector<int64_t> idx(rows, rows + n);                         // extract_rows
  std::vector<T> decay(static_cast<size_t>(n * width));
  const int64_t rc = LookupTable<T>(table, height, width, idx.data(), n, -1, decay.data());
  if (rc != 0) return rc;
  if (mode == 1)
    for (auto& x : decay) x = T((x > T(0)) - (x < T(0)));           // sign
  for (auto& x : decay) x = coeff * x;                              // scale
  for (int64_t i = 0; i < n * width; ++i) out[i] = addend ? addend[i] + decay[i] : decay[i];
  return 0;
}
}  // namespace
extern "C" {
int64_t oracle_sparse_decay_f32(const float* table, int64_t height, int64_t width, const int64_t* rows,
                                int64_t n, float coeff, int mode, const float* addend, float* out) {
  return SparseDecay<float>(table, height, width, rows, n, coeff, mode, addend, out);
}
int64_t oracle_sparse_decay_f64(const double* table, int64_t height, int64_t width, const int64_t* rows,
                                int64_t n, double coeff, int mode, const double* addend, double* out) {
  return SparseDecay<double>(table, height, width, rows, n, coeff, mode, addend, out);
}

// GatherSelectedRows (multi-device reduce): framework/details/reduce_and_gather.h:56-93
// = concatenation in device order.  Implemented by repeated oracle_selected_rows_add_to
// from Python; nothing numeric happens.

int oracle_abi_version() { return 1; }

}  // extern "C"
