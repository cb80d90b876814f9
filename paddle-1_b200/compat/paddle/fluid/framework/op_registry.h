// compat: framework/op_registry.h — kernel registration with the reference's macro names.
// AllOpKernels()[op][{dtype, is_gpu}] holds a std::function that default-constructs a fresh kernel
// object per call and runs Compute (op_registry.h:79-125 of the reference; last registration wins).
// REGISTER_OP_*_KERNEL must be used at global namespace, as in the reference (:171-175).
#pragma once
#include <functional>
#include <map>
#include <string>
#include <typeindex>
#include <utility>
#include "paddle/fluid/framework/operator.h"
#include "paddle/fluid/platform/place.h"

namespace paddle {
namespace framework {

struct OpKernelKey {
  std::type_index dtype;
  bool is_gpu;
  bool operator<(const OpKernelKey& o) const {
    return dtype != o.dtype ? dtype < o.dtype : is_gpu < o.is_gpu;
  }
};
using OpKernelFunc = std::function<void(const ExecutionContext&)>;
using OpKernelMap = std::map<OpKernelKey, OpKernelFunc>;

inline std::map<std::string, OpKernelMap>& AllOpKernels() {
  static std::map<std::string, OpKernelMap> g;
  return g;
}

template <typename PlaceType, size_t I, typename... KernelTypes>
struct OpKernelRegistrarFunctor;

template <typename PlaceType, size_t I, typename... KernelTypes>
struct OpKernelRegistrarFunctor {
  using Tuple = std::tuple<KernelTypes...>;
  void operator()(const char* op_type) const {
    if constexpr (I < sizeof...(KernelTypes)) {
      using KERNEL_TYPE = typename std::tuple_element<I, Tuple>::type;
      using T = typename KERNEL_TYPE::ELEMENT_TYPE;
      OpKernelKey key{std::type_index(typeid(T)), std::is_same<PlaceType, platform::CUDAPlace>::value};
      AllOpKernels()[op_type][key] = [](const ExecutionContext& ctx) { KERNEL_TYPE().Compute(ctx); };
      OpKernelRegistrarFunctor<PlaceType, I + 1, KernelTypes...>()(op_type);
    }
  }
};

template <typename PlaceType, typename... KernelType>
class OpKernelRegistrar {
 public:
  explicit OpKernelRegistrar(const char* op_type) {
    OpKernelRegistrarFunctor<PlaceType, 0, KernelType...>()(op_type);
  }
  void Touch() {}
};

}  // namespace framework
}  // namespace paddle

#define STATIC_ASSERT_GLOBAL_NAMESPACE(uniq_name, msg)                         \
  struct __test_global_namespace_##uniq_name##__ {};                           \
  static_assert(std::is_same< ::__test_global_namespace_##uniq_name##__,       \
                             __test_global_namespace_##uniq_name##__>::value, msg)

#define REGISTER_OP_KERNEL(op_type, DEVICE_TYPE, place_class, ...)                                  \
  STATIC_ASSERT_GLOBAL_NAMESPACE(__reg_op_kernel_##op_type##_##DEVICE_TYPE##__,                     \
                                 "REGISTER_OP_KERNEL must be called in global namespace");          \
  static ::paddle::framework::OpKernelRegistrar<place_class, __VA_ARGS__>                           \
      __op_kernel_registrar_##op_type##_##DEVICE_TYPE##__(#op_type);                                \
  int TouchOpKernelRegistrar_##op_type##_##DEVICE_TYPE() {                                          \
    __op_kernel_registrar_##op_type##_##DEVICE_TYPE##__.Touch();                                    \
    return 0;                                                                                       \
  }

#define REGISTER_OP_CUDA_KERNEL(op_type, ...) \
  REGISTER_OP_KERNEL(op_type, CUDA, ::paddle::platform::CUDAPlace, __VA_ARGS__)
#define REGISTER_OP_CPU_KERNEL(op_type, ...) \
  REGISTER_OP_KERNEL(op_type, CPU, ::paddle::platform::CPUPlace, __VA_ARGS__)
