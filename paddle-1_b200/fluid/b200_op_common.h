// b200_op_common.h — glue shared by the OpKernel wrappers: T -> C-ABI entry point, workspace
// tensor, return-code -> PADDLE_ENFORCE.  Uses only names that exist in the reference tree plus
// include/b200emb.h, so these sources also compile inside a real Paddle checkout.
#pragma once
#include <cstdlib>
#include "b200emb.h"
#include "paddle/fluid/framework/op_registry.h"
#include "paddle/fluid/framework/selected_rows.h"
#include "b200_side_table.h"

namespace paddle {
namespace operators {
namespace b200 {

inline unsigned Flags() {   // FLAGS_b200emb_deterministic in a real tree; env var here
  static const unsigned f = [] {
    const char* e = std::getenv("B200EMB_DETERMINISTIC");
    return (e && e[0] && e[0] != '0') ? B200EMB_DETERMINISTIC : 0u;
  }();
  return f;
}

// temporary from the framework allocator (Tensor::mutable_data -> memory::Alloc in the reference)
inline void* Workspace(framework::Tensor* ws, int64_t n, int64_t d, size_t elem, const platform::Place& place,
                       size_t* bytes) {
  *bytes = b200emb_workspace_bytes(n, d, static_cast<int64_t>(elem));
  return ws->mutable_data<uint8_t>(framework::make_ddim({static_cast<int64_t>(*bytes)}), place);
}

// Workspace + flags for a consumer of a sparse gradient: the plan the forward op started (fused mode) once the
// main stream has been told to wait for it, else a fresh temporary from the framework allocator.
struct UpdateWorkspace {
  framework::Tensor tmp;
  void* ptr = nullptr;
  size_t bytes = 0;
  unsigned flags = 0;
};
inline void PrepareWorkspace(UpdateWorkspace* w, const SparseGradView& g, size_t elem, const platform::Place& place,
                             cudaStream_t stream) {
  w->flags = Flags();
  if (g.plan && !(w->flags & B200EMB_DETERMINISTIC) && g.plan->bytes >= b200emb_workspace_bytes(g.n, g.d, elem)) {
    PADDLE_ENFORCE(cudaStreamWaitEvent(stream, g.plan->done, 0) == cudaSuccess);
    w->ptr = g.plan->ws.data<uint8_t>();
    w->bytes = g.plan->bytes;
    w->flags |= B200EMB_PLAN_READY;
  } else {
    w->ptr = Workspace(&w->tmp, g.n, g.d, elem, place, &w->bytes);
  }
}

// The reference enforces 0 <= id < height on the host (lookup_table_op.h:59-60, sgd_op.h:75-76) or traps the
// device (platform/assert.h:20-29).  The kernels here skip such rows and raise a bit in a per-device status
// word instead; with FLAGS_b200emb_check_ids (env B200EMB_CHECK_IDS=1) every op reads that word back after
// its launch — one stream sync per op, like FLAGS_benchmark (framework/operator.cc:722-724) — and enforces.
inline bool CheckIds() {
  static const bool on = [] {
    const char* e = std::getenv("B200EMB_CHECK_IDS");
    return e && e[0] && e[0] != '0';
  }();
  return on;
}
class StatusWord {
 public:
  StatusWord(const platform::Place& place, cudaStream_t stream) : stream_(stream) {
    if (!CheckIds()) return;
    word_ = t_.mutable_data<int32_t>(framework::make_ddim({1}), place);
    PADDLE_ENFORCE(cudaMemsetAsync(word_, 0, sizeof(int32_t), stream_) == cudaSuccess);
  }
  int32_t* get() { return word_; }
  void Enforce(const char* op) {
    if (!word_) return;
    int32_t h = 0;
    PADDLE_ENFORCE(cudaMemcpyAsync(&h, word_, sizeof(h), cudaMemcpyDeviceToHost, stream_) == cudaSuccess);
    PADDLE_ENFORCE(cudaStreamSynchronize(stream_) == cudaSuccess);
    PADDLE_ENFORCE_EQ(h, 0, "%s: an id / row is outside [0, height) (status %d)", op, h);
  }

 private:
  framework::Tensor t_;
  int32_t* word_ = nullptr;
  cudaStream_t stream_;
};

#define B200EMB_CHECK(call)                                                              \
  do {                                                                                   \
    int rc__ = (call);                                                                   \
    PADDLE_ENFORCE_EQ(rc__, 0, #call " failed (cuda error %d)", b200emb_last_cuda_error()); \
  } while (0)

template <typename T> struct Api;
template <> struct Api<float> {
  static constexpr auto lookup = b200emb_lookup_table_f32;
  static constexpr auto grad_sparse = b200emb_lookup_table_grad_sparse_f32;
  static constexpr auto add_to_tensor = b200emb_selected_rows_add_to_tensor_f32;
  static constexpr auto merge_add = b200emb_merge_add_f32;
  static constexpr auto sgd = b200emb_sparse_sgd_f32;
  static constexpr auto adam = b200emb_sparse_adam_f32;
};
template <> struct Api<double> {
  static constexpr auto lookup = b200emb_lookup_table_f64;
  static constexpr auto grad_sparse = b200emb_lookup_table_grad_sparse_f64;
  static constexpr auto add_to_tensor = b200emb_selected_rows_add_to_tensor_f64;
  static constexpr auto merge_add = b200emb_merge_add_f64;
  static constexpr auto sgd = b200emb_sparse_sgd_f64;
  static constexpr auto adam = b200emb_sparse_adam_f64;
};

}  // namespace b200
}  // namespace operators
}  // namespace paddle
