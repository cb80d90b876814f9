// compat: platform/enforce.h — PADDLE_ENFORCE family over a std::exception (EnforceNotMet), enough
// for the op kernels of the sparse-embedding path.  Mirrors the names/semantics of the reference's
// platform/enforce.h:69-131,244-281 (no stack trace, printf-style message kept).
#pragma once
#include <cstdio>
#include <exception>
#include <sstream>
#include <string>
#include "paddle/fluid/platform/hostdevice.h"

namespace paddle {
namespace platform {

struct EnforceNotMet : public std::exception {
  std::string msg;
  EnforceNotMet(const std::string& m, const char* file, int line) {
    std::ostringstream os;
    os << m << " at [" << file << ":" << line << "]";
    msg = os.str();
  }
  const char* what() const noexcept override { return msg.c_str(); }
};

namespace detail {
inline std::string Format() { return std::string(); }
inline std::string Format(const char* s) { return std::string(s); }
inline std::string Format(const std::string& s) { return s; }
template <typename... Args>
std::string Format(const char* fmt, Args... args) {
  char buf[512];
  std::snprintf(buf, sizeof(buf), fmt, args...);
  return std::string(buf);
}
template <typename T>
bool Truthy(const T& v) { return static_cast<bool>(v); }
}  // namespace detail

}  // namespace platform
}  // namespace paddle

#define PADDLE_THROW(...) \
  throw ::paddle::platform::EnforceNotMet(::paddle::platform::detail::Format(__VA_ARGS__), __FILE__, __LINE__)

#define PADDLE_ENFORCE(cond, ...)                                                              \
  do {                                                                                         \
    if (!::paddle::platform::detail::Truthy(cond))                                             \
      throw ::paddle::platform::EnforceNotMet(                                                 \
          std::string("Enforce failed: " #cond " ") + ::paddle::platform::detail::Format(__VA_ARGS__), \
          __FILE__, __LINE__);                                                                 \
  } while (0)

#define PADDLE_ENFORCE_NOT_NULL(p, ...) PADDLE_ENFORCE((p) != nullptr, ##__VA_ARGS__)

#define B200_COMPAT_BINARY_ENFORCE(a, b, op, ...)                                              \
  do {                                                                                         \
    if (!((a)op(b))) {                                                                         \
      std::ostringstream os__;                                                                 \
      os__ << "Enforce failed. Expected " #a " " #op " " #b ", but received " << (a) << " vs " << (b) << ". " \
           << ::paddle::platform::detail::Format(__VA_ARGS__);                                 \
      throw ::paddle::platform::EnforceNotMet(os__.str(), __FILE__, __LINE__);                 \
    }                                                                                          \
  } while (0)

#define PADDLE_ENFORCE_EQ(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, ==, ##__VA_ARGS__)
#define PADDLE_ENFORCE_NE(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, !=, ##__VA_ARGS__)
#define PADDLE_ENFORCE_GT(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, >, ##__VA_ARGS__)
#define PADDLE_ENFORCE_GE(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, >=, ##__VA_ARGS__)
#define PADDLE_ENFORCE_LT(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, <, ##__VA_ARGS__)
#define PADDLE_ENFORCE_LE(a, b, ...) B200_COMPAT_BINARY_ENFORCE(a, b, <=, ##__VA_ARGS__)

// glog's VLOG(n) << ... : swallowed.
#ifndef VLOG
namespace paddle { namespace platform { struct NullLog { template <typename T> NullLog& operator<<(const T&) { return *this; } }; } }
#define VLOG(n) if (false) ::paddle::platform::NullLog()
#endif
