// compat: framework/details/op_handle_base.h — the operator nodes of ParallelExecutor's SSA graph
// (op_handle_base.h:32-110 of the reference): Name(), Run(use_cuda) -> RunImpl(), inputs / outputs as
// VarHandleBase*, one DeviceContext per place.  Event bookkeeping between op handles (RunAndRecordEvent,
// WaitInputVarGenerated) is reduced to stream order: every handle here enqueues on the places' own streams.
#pragma once
#include <map>
#include <string>
#include <vector>
#include "paddle/fluid/framework/details/var_handle.h"
#include "paddle/fluid/platform/device_context.h"

namespace paddle {
namespace framework {
namespace details {

class OpHandleBase {
 public:
  virtual ~OpHandleBase() {}
  virtual std::string Name() const = 0;
  void Run(bool /*use_cuda*/) { RunImpl(); }
  void AddInput(VarHandleBase* in) {
    inputs_.push_back(in);
    in->pending_ops_.insert(this);
  }
  void AddOutput(VarHandleBase* out) {
    outputs_.push_back(out);
    out->generated_op_ = this;
  }
  virtual bool IsMultiDeviceTransfer() { return false; }
  const platform::DeviceContext* DeviceContext(platform::Place place) { return dev_ctxes_[place]; }
  void SetDeviceContext(platform::Place place, platform::DeviceContext* ctx) { dev_ctxes_[place] = ctx; }
  const std::vector<VarHandleBase*>& Inputs() const { return inputs_; }
  const std::vector<VarHandleBase*>& Outputs() const { return outputs_; }

 protected:
  virtual void RunImpl() = 0;
  std::vector<VarHandleBase*> inputs_;
  std::vector<VarHandleBase*> outputs_;
  std::map<platform::Place, platform::DeviceContext*> dev_ctxes_;
};

}  // namespace details
}  // namespace framework
}  // namespace paddle
