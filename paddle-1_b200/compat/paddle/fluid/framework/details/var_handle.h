// compat: framework/details/var_handle.h — the variable nodes of ParallelExecutor's SSA graph, reduced to what
// an op handle reads: a VarHandle names one version of one runtime variable in one local scope on one place
// (var_handle.h:87-113 of the reference).  The ir::Node plumbing of the reference's graph is not mirrored.
#pragma once
#include <string>
#include <unordered_set>
#include <utility>
#include "paddle/fluid/platform/place.h"

namespace paddle {
namespace framework {
namespace details {

class OpHandleBase;

struct VarHandleBase {
  virtual ~VarHandleBase() {}
  virtual std::string DebugString() const = 0;
  OpHandleBase* GeneratedOp() { return generated_op_; }
  const std::unordered_set<OpHandleBase*>& PendingOps() const { return pending_ops_; }
  OpHandleBase* generated_op_{nullptr};            // the operator that produces this version (nullptr: a root)
  std::unordered_set<OpHandleBase*> pending_ops_;  // operators waiting for it
};

struct VarHandle : public VarHandleBase {
  VarHandle(size_t version, size_t scope_index, std::string name, platform::Place place)
      : version_(version), scope_idx_(scope_index), name_(std::move(name)), place_(std::move(place)) {}
  std::string DebugString() const override { return name_ + "@scope" + std::to_string(scope_idx_); }
  size_t version_;
  size_t scope_idx_;
  std::string name_;
  platform::Place place_;
};

struct DummyVarHandle : public VarHandleBase {
  std::string DebugString() const override { return "dummy"; }
};

}  // namespace details
}  // namespace framework
}  // namespace paddle
