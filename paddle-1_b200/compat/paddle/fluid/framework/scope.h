// compat: framework/scope.h — name -> Variable map (Var creates, FindVar looks up), scope.h:39-80.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include "paddle/fluid/framework/variable.h"

namespace paddle {
namespace framework {
class Scope {
 public:
  Variable* Var(const std::string& name) {
    auto it = vars_.find(name);
    if (it != vars_.end()) return it->second.get();
    auto* v = new Variable();
    vars_[name].reset(v);
    return v;
  }
  Variable* FindVar(const std::string& name) const {
    auto it = vars_.find(name);
    return it == vars_.end() ? nullptr : it->second.get();
  }
  void EraseVar(const std::string& name) { vars_.erase(name); }

 private:
  std::unordered_map<std::string, std::unique_ptr<Variable>> vars_;
};
}  // namespace framework
}  // namespace paddle
