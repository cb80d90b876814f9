// compat: framework/operator.h — the part of the op/kernel interface a kernel sees:
// OperatorBase (type, slot->variable-name maps, attributes), ExecutionContext (Input/Output/
// InputVar/OutputVar/Multi*, Attr<T>, GetPlace, device contexts) and OpKernel<T>.  Semantics follow
// the reference's operator.h:154-311 and operator.cc:350-420: Input<Tensor>/Input<LoDTensor> on a
// SelectedRows variable yields its value().
#pragma once
#include <map>
#include <string>
#include <type_traits>
#include <variant>
#include <vector>
#include "paddle/fluid/framework/lod_tensor.h"
#include "paddle/fluid/framework/scope.h"
#include "paddle/fluid/framework/selected_rows.h"
#include "paddle/fluid/platform/device_context.h"

namespace paddle {
namespace framework {

constexpr char kGradVarSuffix[] = "@GRAD";
inline std::string GradVarName(const std::string& var_name) { return var_name + kGradVarSuffix; }

using Attribute = std::variant<bool, int, int64_t, float, std::string, std::vector<int>>;
using AttributeMap = std::map<std::string, Attribute>;
using VariableNameMap = std::map<std::string, std::vector<std::string>>;

class OperatorBase {
 public:
  OperatorBase(const std::string& type, const VariableNameMap& inputs, const VariableNameMap& outputs,
               const AttributeMap& attrs)
      : type_(type), inputs_(inputs), outputs_(outputs), attrs_(attrs) {}
  const std::string& Type() const { return type_; }
  const VariableNameMap& Inputs() const { return inputs_; }
  const VariableNameMap& Outputs() const { return outputs_; }
  const std::vector<std::string>& Inputs(const std::string& name) const {
    static const std::vector<std::string> kEmpty;
    auto it = inputs_.find(name);
    return it == inputs_.end() ? kEmpty : it->second;
  }
  const std::vector<std::string>& Outputs(const std::string& name) const {
    static const std::vector<std::string> kEmpty;
    auto it = outputs_.find(name);
    return it == outputs_.end() ? kEmpty : it->second;
  }
  std::string Input(const std::string& name) const {
    auto& v = Inputs(name);
    PADDLE_ENFORCE_LE(v.size(), 1UL, "Operator %s's input %s should contain only one variable.", type_.c_str(), name.c_str());
    return v.empty() ? std::string() : v[0];
  }
  std::string Output(const std::string& name) const {
    auto& v = Outputs(name);
    PADDLE_ENFORCE_LE(v.size(), 1UL, "Operator %s's output %s should contain only one variable.", type_.c_str(), name.c_str());
    return v.empty() ? std::string() : v[0];
  }
  template <typename T>
  const T& Attr(const std::string& name) const {
    auto it = attrs_.find(name);
    PADDLE_ENFORCE(it != attrs_.end(), "%s should be in AttributeMap", name.c_str());
    PADDLE_ENFORCE(std::holds_alternative<T>(it->second), "attribute %s has another type", name.c_str());
    return std::get<T>(it->second);
  }
  const AttributeMap& Attrs() const { return attrs_; }

 private:
  std::string type_;
  VariableNameMap inputs_, outputs_;
  AttributeMap attrs_;
};

class ExecutionContext {
 public:
  ExecutionContext(const OperatorBase& op, const Scope& scope, const platform::DeviceContext& device_context)
      : op_(op), scope_(scope), device_context_(device_context) {}
  const OperatorBase& op() const { return op_; }
  const Scope& scope() const { return scope_; }
  template <typename T>
  inline const T& Attr(const std::string& name) const { return op_.Attr<T>(name); }
  size_t InputSize(const std::string& name) const { return op_.Inputs(name).size(); }
  size_t OutputSize(const std::string& name) const { return op_.Outputs(name).size(); }

  const Variable* InputVar(const std::string& name) const {
    auto ipt = op_.Input(name);
    return ipt.empty() ? nullptr : scope_.FindVar(ipt);
  }
  Variable* OutputVar(const std::string& name) const {
    auto opt = op_.Output(name);
    return opt.empty() ? nullptr : scope_.FindVar(opt);
  }
  const std::vector<const Variable*> MultiInputVar(const std::string& name) const {
    std::vector<const Variable*> res;
    for (auto& n : op_.Inputs(name)) res.push_back(n.empty() ? nullptr : scope_.FindVar(n));
    return res;
  }
  std::vector<Variable*> MultiOutputVar(const std::string& name) const {
    std::vector<Variable*> res;
    for (auto& n : op_.Outputs(name)) res.push_back(n.empty() ? nullptr : scope_.FindVar(n));
    return res;
  }

  template <typename T>
  const T* Input(const std::string& name) const {
    auto* var = InputVar(name);
    return var == nullptr ? nullptr : ConstCast<T>(var);
  }
  template <typename T>
  T* Output(const std::string& name) const {
    auto* var = OutputVar(name);
    return var == nullptr ? nullptr : MutableCast<T>(var);
  }
  template <typename T>
  const std::vector<const T*> MultiInput(const std::string& name) const {
    std::vector<const T*> res;
    for (auto* v : MultiInputVar(name)) res.push_back(v == nullptr ? nullptr : ConstCast<T>(v));
    return res;
  }
  template <typename T>
  std::vector<T*> MultiOutput(const std::string& name) const {
    std::vector<T*> res;
    for (auto* v : MultiOutputVar(name)) res.push_back(v == nullptr ? nullptr : MutableCast<T>(v));
    return res;
  }

  platform::Place GetPlace() const { return device_context_.GetPlace(); }
  template <typename DeviceContextType>
  const DeviceContextType& device_context() const {
    return *reinterpret_cast<const DeviceContextType*>(&device_context_);
  }
  const platform::DeviceContext& device_context() const { return device_context_; }
#ifdef PADDLE_WITH_CUDA
  const inline platform::CUDADeviceContext& cuda_device_context() const {
    PADDLE_ENFORCE(platform::is_gpu_place(device_context_.GetPlace()));
    return *reinterpret_cast<const platform::CUDADeviceContext*>(&device_context_);
  }
#endif

 private:
  // Tensor / LoDTensor requested from a SelectedRows variable -> its value (operator.cc:350-359)
  template <typename T>
  static const T* ConstCast(const Variable* var) {
    if constexpr (std::is_same<T, Tensor>::value || std::is_same<T, LoDTensor>::value) {
      if (var->IsType<LoDTensor>()) return static_cast<const T*>(&var->Get<LoDTensor>());
      if (var->IsType<SelectedRows>()) return static_cast<const T*>(&var->Get<SelectedRows>().value());
      PADDLE_THROW("Variable type must be LoDTensor/SelectedRows.");
    } else {
      return &var->Get<T>();
    }
  }
  template <typename T>
  static T* MutableCast(Variable* var) {
    if constexpr (std::is_same<T, Tensor>::value || std::is_same<T, LoDTensor>::value) {
      if (var->IsType<SelectedRows>()) return static_cast<T*>(var->GetMutable<SelectedRows>()->mutable_value());
      return static_cast<T*>(var->GetMutable<LoDTensor>());
    } else {
      return var->GetMutable<T>();
    }
  }
  const OperatorBase& op_;
  const Scope& scope_;
  const platform::DeviceContext& device_context_;
};

class OpKernelBase {
 public:
  virtual void Compute(const ExecutionContext& context) const = 0;
  virtual ~OpKernelBase() = default;
};

template <typename T>
class OpKernel : public OpKernelBase {
 public:
  using ELEMENT_TYPE = T;
};

}  // namespace framework
}  // namespace paddle
