// regularize.cu — the sparse weight-decay chain and split_selected_rows, sm_100a.
//
// k_sparse_decay: for a SelectedRows gradient the reference's regularizers emit four ops
// (python/paddle/fluid/regularizer.py:146-170 L2, :226-253 L1, then :64-69):
//     extract_rows(grad) -> idx            operators/extract_rows_op.cc:49-77  (a copy of the row ids)
//     lookup_table(W, idx) -> decay        operators/lookup_table_op.h:38-64   (one row per occurrence)
//     [sign(decay)]                        operators/sign_op.h:34              (L1 only)
//     scale(decay, coeff)                  operators/scale_op.h:38-56          (scale * x)
//     elementwise_add(grad, decay)         (the caller in append_regularization_ops)
// i.e. out[i,:] = grad[i,:] + coeff * f(W[rows[i],:]) with every product and sum rounded on its own.
// One kernel does all of it: each gradient row and each table row is read once, the regularised
// gradient written once; the decay tensor never exists.  With addend == nullptr the kernel
// produces the decay term alone (the first four ops).
//
// split_selected_rows (operators/split_selected_rows_op.h:24-102): range split of a SelectedRows by
// `height_sections`; k_section_of maps every row to its section (FindOutIdx, :24-32), the usual
// stable sort plan orders the positions by section, k_rebase_rows writes rows - section start.
#include <algorithm>

#include "common.cuh"
#include "internal.h"

namespace b200emb {
namespace {

template <typename T, int LPR, int NV, bool VEC, int MODE>
__global__ void __launch_bounds__(kBlock, 4)
k_sparse_decay(const T* __restrict__ table, uint32_t height, int d, const int64_t* __restrict__ rows,
               int n, T coeff, const T* __restrict__ addend, T* __restrict__ out, int32_t* status) {
  constexpr int EPV = Epv<T, VEC>::value;
  constexpr int kGroups = kBlock / LPR;
  constexpr int kRows = 4;
  using A = Arith<T>;
  const int l = threadIdx.x % LPR;
  const int group = blockIdx.x * kGroups + threadIdx.x / LPR;
  const int num_groups = gridDim.x * kGroups;
  bool bad = false;
  for (int base = group; base < n; base += num_groups * kRows) {
    int64_t from[kRows];
    bool live[kRows], ok[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      const int e = base + r * num_groups;
      live[r] = e < n;
      const int64_t id = live[r] ? rows[e] : 0;
      ok[r] = live[r] && id >= 0 && id < static_cast<int64_t>(height);
      bad |= live[r] && !ok[r];
      from[r] = ok[r] ? id : 0;
    }
    for (int col0 = 0; col0 < d; col0 += LPR * NV * EPV) {
      T w[kRows][NV][EPV], g[kRows][NV][EPV];
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          const bool in = live[r] && col < d;
          if (in && ok[r]) load_vec<T, VEC>(table + from[r] * d + col, w[r][k]);
          else {
#pragma unroll
            for (int e = 0; e < EPV; ++e) w[r][k][e] = T(0);
          }
          if (in && addend)
            load_vec_stream<T, VEC>(addend + static_cast<int64_t>(base + r * num_groups) * d + col, g[r][k]);
        }
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int col = col0 + (l + k * LPR) * EPV;
          if (!(live[r] && col < d)) continue;
          T o[EPV];
#pragma unroll
          for (int e = 0; e < EPV; ++e) {
            T x = w[r][k][e];
            if (MODE == 1) x = static_cast<T>((T(0) < x) - (x < T(0)));   // Eigen's sign(): +-1, 0 for 0 / NaN
            const T dec = ok[r] ? A::mul(coeff, x) : T(0);
            o[e] = addend ? A::add(g[r][k][e], dec) : dec;
          }
          store_vec_stream<T, VEC>(out + static_cast<int64_t>(base + r * num_groups) * d + col, o);
        }
    }
  }
  if (bad && status) atomicOr(status, 1);
}

constexpr int kMaxSections = 64;
struct Sections {
  int64_t abs[kMaxSections];   // absolute start row of every section
  int count;
};

// FindOutIdx (split_selected_rows_op.h:24-32): the first section whose successor starts above the row;
// rows past the last start — and nothing else — fall into the last section, exactly as there.
__global__ void __launch_bounds__(256)
k_section_of(const int64_t* __restrict__ rows, int n, Sections sec, int64_t* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int64_t r = rows[i];
    int s = sec.count - 1;
    for (int j = 1; j < sec.count; ++j)
      if (r < sec.abs[j]) { s = j - 1; break; }
    out[i] = s;
  }
}

__global__ void __launch_bounds__(256)
k_rebase_rows(const int64_t* __restrict__ rows, const uint32_t* __restrict__ skey,
              const uint32_t* __restrict__ perm, int n, Sections sec, int64_t* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    out[i] = rows[perm[i]] - sec.abs[skey[i]];
}

}  // namespace

template <typename T>
cudaError_t launch_sparse_decay(const T* table, int64_t height, int64_t d, const int64_t* rows, int64_t n,
                                T coeff, int mode, const T* addend, T* out, int32_t* status,
                                cudaStream_t s) {
  const bool aligned = is_aligned16(table) && is_aligned16(out) && (!addend || is_aligned16(addend));
  cudaError_t err = cudaSuccess;
  dispatch_row_shape<T>(d, aligned, [&]<int LPR, int NV, bool VEC>() {
    constexpr int kGroups = kBlock / LPR;
    const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, kGroups * 4), 148 * 4));
    const auto h = static_cast<uint32_t>(height);
    const int ni = static_cast<int>(n), di = static_cast<int>(d);
    if (mode == 1)
      k_sparse_decay<T, LPR, NV, VEC, 1><<<grid, kBlock, 0, s>>>(table, h, di, rows, ni, coeff, addend, out, status);
    else
      k_sparse_decay<T, LPR, NV, VEC, 0><<<grid, kBlock, 0, s>>>(table, h, di, rows, ni, coeff, addend, out, status);
    err = cudaGetLastError();
  });
  return err;
}

template cudaError_t launch_sparse_decay<float>(const float*, int64_t, int64_t, const int64_t*, int64_t, float,
                                                int, const float*, float*, int32_t*, cudaStream_t);
template cudaError_t launch_sparse_decay<double>(const double*, int64_t, int64_t, const int64_t*, int64_t,
                                                 double, int, const double*, double*, int32_t*, cudaStream_t);

int max_sections() { return kMaxSections; }

cudaError_t launch_section_of(const int64_t* rows, int64_t n, const int64_t* height_sections, int n_sections,
                              int64_t* out, cudaStream_t s) {
  Sections sec{};
  sec.count = n_sections;
  for (int j = 1; j < n_sections; ++j) sec.abs[j] = sec.abs[j - 1] + height_sections[j - 1];
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, 256), 148 * 8));
  k_section_of<<<grid, 256, 0, s>>>(rows, static_cast<int>(n), sec, out);
  return cudaGetLastError();
}

cudaError_t launch_rebase_rows(const int64_t* rows, const SortedPlan& plan, int64_t n,
                               const int64_t* height_sections, int n_sections, int64_t* out,
                               cudaStream_t s) {
  Sections sec{};
  sec.count = n_sections;
  for (int j = 1; j < n_sections; ++j) sec.abs[j] = sec.abs[j - 1] + height_sections[j - 1];
  const int grid = static_cast<int>(std::min<int64_t>(ceil_div(n, 256), 148 * 8));
  k_rebase_rows<<<grid, 256, 0, s>>>(rows, plan.keys, plan.perm, static_cast<int>(n), sec, out);
  return cudaGetLastError();
}

}  // namespace b200emb
