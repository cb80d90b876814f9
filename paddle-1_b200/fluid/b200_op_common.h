// b200_op_common.h — glue shared by the OpKernel wrappers: T -> C-ABI entry point, workspace
// tensor, return-code -> PADDLE_ENFORCE.  Uses only names that exist in the reference tree plus
// include/b200emb.h, so these sources also compile inside a real Paddle checkout.
#pragma once
#include <cstdlib>
#include "b200emb.h"
#include "paddle/fluid/framework/op_registry.h"
#include "paddle/fluid/framework/selected_rows.h"

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
