// compat: operators/detail/safe_ref.h — Ref(ptr, msg...) = checked dereference.
#pragma once
#include "paddle/fluid/platform/enforce.h"
namespace paddle {
namespace operators {
namespace detail {
template <typename T, typename... ARGS>
inline T& Ref(T* ptr, ARGS&&... args) {
  PADDLE_ENFORCE(ptr != nullptr, args...);
  return *ptr;
}
}  // namespace detail
}  // namespace operators
}  // namespace paddle
