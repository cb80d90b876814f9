// ref_driver.cc — drives the REFERENCE'S OWN, UNMODIFIED CPU kernels for the sparse-embedding path.
//
// TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/oracle.cc's header for who may use oracle/).
// The reference cannot be built with its own build system here (Boost, Eigen, glog, gflags,
// protobuf are cmake ExternalProject downloads).  Its kernels for this path, however, are plain
// C++ against a small slice of the Fluid API, so this translation unit includes the reference's
// sources where they lie under /root/reference —
//     paddle/fluid/operators/lookup_table_op.h, sgd_op.h, adam_op.h, sum_op.h,
//     paddle/fluid/operators/math/selected_rows_functor.{h,cc}   (the .cc is compiled alongside)
// — with the repo's own compat/ headers (same names, own code) first on the include path, and
// exposes each kernel behind a C function.  Nothing from the reference is copied into the repo;
// the output goes to oracle/_ref/ (git-ignored, shipped to the GPU box as a built artefact).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "paddle/fluid/operators/adam_op.h"
#include "paddle/fluid/operators/lookup_table_op.h"
#include "paddle/fluid/operators/math/selected_rows_functor.h"
#include "paddle/fluid/operators/sgd_op.h"
#include "paddle/fluid/operators/sum_op.h"

namespace fw = paddle::framework;
namespace pf = paddle::platform;
namespace ops = paddle::operators;

namespace {

thread_local std::string g_last_error;

template <typename T>
fw::LoDTensor* PutTensor(fw::Scope* scope, const std::string& name, T* data, std::vector<int64_t> dims) {
  auto* t = scope->Var(name)->GetMutable<fw::LoDTensor>();
  t->ShareExternalData(data, fw::make_ddim(dims), pf::CPUPlace());
  return t;
}

template <typename T>
fw::SelectedRows* PutSelectedRows(fw::Scope* scope, const std::string& name, const int64_t* rows, int64_t n,
                                  T* value, int64_t d, int64_t height) {
  auto* sr = scope->Var(name)->GetMutable<fw::SelectedRows>();
  sr->set_rows(std::vector<int64_t>(rows, rows + n));
  sr->set_height(height);
  sr->mutable_value()->ShareExternalData(value, fw::make_ddim({n, d}), pf::CPUPlace());
  return sr;
}

template <typename Kernel>
int RunKernel(const std::string& type, const fw::VariableNameMap& in, const fw::VariableNameMap& out,
              const fw::AttributeMap& attrs, fw::Scope* scope) {
  try {
    fw::OperatorBase op(type, in, out, attrs);
    pf::CPUDeviceContext dev;
    fw::ExecutionContext ctx(op, *scope, dev);
    Kernel().Compute(ctx);
    return 0;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return -1;
  }
}

}  // namespace

extern "C" {

const char* ref_last_error() { return g_last_error.c_str(); }
int ref_abi_version() { return 1; }

// LookupTableKernel<T>::Compute (operators/lookup_table_op.h:36-84)
#define REF_DEFINE(T, S)                                                                              \
  int ref_lookup_table_##S(const T* w, int64_t v, int64_t d, const int64_t* ids, int64_t n,           \
                           int64_t padding_idx, T* out) {                                             \
    fw::Scope scope;                                                                                  \
    PutTensor(&scope, "W", w, {v, d});                                                                \
    PutTensor(&scope, "Ids", ids, {n, 1});                                                            \
    PutTensor(&scope, "Out", out, {n, d});                                                            \
    return RunKernel<ops::LookupTableKernel<T>>("lookup_table", {{"W", {"W"}}, {"Ids", {"Ids"}}},     \
                                                {{"Out", {"Out"}}},                                   \
                                                {{"padding_idx", padding_idx}, {"is_sparse", true}},  \
                                                &scope);                                              \
  }                                                                                                   \
  /* LookupTableGradKernel<T>, sparse branch (lookup_table_op.h:106-134) */                           \
  int ref_lookup_table_grad_sparse_##S(const int64_t* ids, int64_t n, const T* dout, int64_t d,       \
                                       int64_t height, int64_t* rows_out, T* value_out) {             \
    fw::Scope scope;                                                                                  \
    std::vector<T> dummy_w(1);                                                                        \
    auto* w = scope.Var("W")->GetMutable<fw::LoDTensor>();                                            \
    w->ShareExternalData(dummy_w.data(), fw::make_ddim({height, d}), pf::CPUPlace());                 \
    PutTensor(&scope, "Ids", ids, {n, 1});                                                            \
    PutTensor(&scope, "Out@GRAD", dout, {n, d});                                                      \
    auto* g = scope.Var("W@GRAD")->GetMutable<fw::SelectedRows>();                                    \
    int rc = RunKernel<ops::LookupTableGradKernel<T>>(                                                \
        "lookup_table_grad", {{"W", {"W"}}, {"Ids", {"Ids"}}, {"Out@GRAD", {"Out@GRAD"}}},            \
        {{"W@GRAD", {"W@GRAD"}}}, {{"is_sparse", true}, {"padding_idx", int64_t(-1)}}, &scope);       \
    if (rc != 0) return rc;                                                                           \
    if (g->height() != height || static_cast<int64_t>(g->rows().size()) != n) return -2;              \
    for (int64_t i = 0; i < n; ++i) rows_out[i] = g->rows()[i];                                       \
    std::memcpy(value_out, g->value().data<T>(), sizeof(T) * n * d);                                  \
    return 0;                                                                                         \
  }                                                                                                   \
  /* SumKernel<CPU,T>, SelectedRows output (sum_op.h:71-145): k inputs concatenated */                \
  int64_t ref_sum_selected_rows_##S(int k, const int64_t* const* rows, const int64_t* ns,             \
                                    const T* const* values, int64_t d, int64_t height,                \
                                    int64_t* out_rows, T* out_value) {                                \
    fw::Scope scope;                                                                                  \
    std::vector<std::string> names;                                                                   \
    for (int i = 0; i < k; ++i) {                                                                     \
      names.push_back("X" + std::to_string(i));                                                       \
      PutSelectedRows(&scope, names.back(), rows[i], ns[i], const_cast<T*>(values[i]), d, height);    \
    }                                                                                                 \
    auto* out = scope.Var("Out")->GetMutable<fw::SelectedRows>();                                     \
    out->set_height(height);                                                                          \
    int rc = RunKernel<ops::SumKernel<pf::CPUDeviceContext, T>>("sum", {{"X", names}},                \
                                                               {{"Out", {"Out"}}}, {}, &scope);       \
    if (rc != 0) return rc;                                                                           \
    const int64_t total = static_cast<int64_t>(out->rows().size());                                   \
    for (int64_t i = 0; i < total; ++i) out_rows[i] = out->rows()[i];                                 \
    if (total) std::memcpy(out_value, out->value().data<T>(), sizeof(T) * total * d);                 \
    return total;                                                                                     \
  }                                                                                                   \
  /* math::SelectedRowsAddToTensor<CPU,T> (selected_rows_functor.cc:153-178) */                       \
  int ref_selected_rows_add_to_tensor_##S(const int64_t* rows, int64_t n, const T* value, int64_t d,  \
                                          int64_t height, T* tensor) {                                \
    try {                                                                                             \
      fw::SelectedRows in(std::vector<int64_t>(rows, rows + n), height);                              \
      in.mutable_value()->ShareExternalData(value, fw::make_ddim({n, d}), pf::CPUPlace());            \
      fw::Tensor t;                                                                                   \
      t.ShareExternalData(tensor, fw::make_ddim({height, d}), pf::CPUPlace());                        \
      pf::CPUDeviceContext dev;                                                                       \
      ops::math::SelectedRowsAddToTensor<pf::CPUDeviceContext, T>()(dev, in, &t);                     \
      return 0;                                                                                       \
    } catch (const std::exception& e) { g_last_error = e.what(); return -1; }                         \
  }                                                                                                   \
  /* math::scatter::MergeAdd<CPU,T> (selected_rows_functor.cc:193-228) */                             \
  int64_t ref_merge_add_##S(const int64_t* rows, int64_t n, const T* value, int64_t d,                \
                            int64_t height, int64_t* out_rows, T* out_value) {                        \
    try {                                                                                             \
      fw::SelectedRows in(std::vector<int64_t>(rows, rows + n), height);                              \
      in.mutable_value()->ShareExternalData(value, fw::make_ddim({n, d}), pf::CPUPlace());            \
      pf::CPUDeviceContext dev;                                                                       \
      fw::SelectedRows merged = ops::math::scatter::MergeAdd<pf::CPUDeviceContext, T>()(dev, in);     \
      const int64_t u = static_cast<int64_t>(merged.rows().size());                                   \
      for (int64_t i = 0; i < u; ++i) out_rows[i] = merged.rows()[i];                                 \
      std::memcpy(out_value, merged.value().data<T>(), sizeof(T) * u * d);                            \
      return u;                                                                                       \
    } catch (const std::exception& e) { g_last_error = e.what(); return -1; }                         \
  }                                                                                                   \
  /* SGDOpKernel<T>, SelectedRows-gradient branch (sgd_op.h:47-81) */                                 \
  int ref_sgd_sparse_##S(T* param, int64_t v, int64_t d, const int64_t* rows, int64_t n,              \
                         const T* grad, const T* lr) {                                                \
    fw::Scope scope;                                                                                  \
    PutTensor(&scope, "Param", param, {v, d});                                                        \
    PutSelectedRows(&scope, "Grad", rows, n, const_cast<T*>(grad), d, v);                             \
    PutTensor(&scope, "LearningRate", lr, {1});                                                       \
    return RunKernel<ops::SGDOpKernel<T>>(                                                            \
        "sgd", {{"Param", {"Param"}}, {"Grad", {"Grad"}}, {"LearningRate", {"LearningRate"}}},        \
        {{"ParamOut", {"Param"}}}, {}, &scope);                                                       \
  }                                                                                                   \
  /* AdamOpKernel<CPU,T>, SelectedRows branch = MergeAdd + SparseAdamFunctor (adam_op.h:282-321) */   \
  int ref_adam_sparse_##S(T* param, T* m1, T* m2, int64_t v, int64_t d, const int64_t* rows,          \
                          int64_t n, const T* grad, const T* lr, const T* b1p, const T* b2p,          \
                          float beta1, float beta2, float epsilon) {                                  \
    fw::Scope scope;                                                                                  \
    PutTensor(&scope, "Param", param, {v, d});                                                        \
    PutTensor(&scope, "Moment1", m1, {v, d});                                                         \
    PutTensor(&scope, "Moment2", m2, {v, d});                                                         \
    PutSelectedRows(&scope, "Grad", rows, n, const_cast<T*>(grad), d, v);                             \
    PutTensor(&scope, "LearningRate", lr, {1});                                                       \
    PutTensor(&scope, "Beta1Pow", b1p, {1});                                                          \
    PutTensor(&scope, "Beta2Pow", b2p, {1});                                                          \
    return RunKernel<ops::AdamOpKernel<pf::CPUDeviceContext, T>>(                                     \
        "adam",                                                                                       \
        {{"Param", {"Param"}}, {"Grad", {"Grad"}}, {"LearningRate", {"LearningRate"}},                \
         {"Moment1", {"Moment1"}}, {"Moment2", {"Moment2"}}, {"Beta1Pow", {"Beta1Pow"}},              \
         {"Beta2Pow", {"Beta2Pow"}}},                                                                 \
        {{"ParamOut", {"Param"}}, {"Moment1Out", {"Moment1"}}, {"Moment2Out", {"Moment2"}}},          \
        {{"beta1", beta1}, {"beta2", beta2}, {"epsilon", epsilon}}, &scope);                          \
  }

REF_DEFINE(float, f32)
REF_DEFINE(double, f64)

}  // extern "C"
