// sparse_all_to_all_op_handle.h — the op handles multi_devices_graph_pass emits for a row-sharded embedding
// table, next to framework/details/reduce_op_handle.h / broadcast_op_handle.h in a Paddle tree.
//
// The reference decides per gradient (framework/details/multi_devices_graph_pass.cc:411-426): for a sparse one
// (IsSparseGradient, :469-475) it emits CreateReduceOp (:660-703) — ReduceOpHandle::RunImpl gathers every
// device's SelectedRows on one device (reduce_op_handle.cc:88-94, reduce_and_gather.h:56-93) — lets the optimizer
// op run there, and CreateBroadcastOp (:490) copies the whole parameter back to every replica
// (broadcast_op_handle.cc:71-122).  With the table sharded by row over the devices (owner = id % G) the pass
// emits instead, per embedding parameter:
//   ShardedLookupOpHandle     in place of the G lookup_table ops:   Out_d = table[Ids_d]  on every device d
//   SparseAllToAllOpHandle    in place of Reduce + sgd|adam + Broadcast: every device's sparse gradient
//                             {rows = Ids_d, value = dOut_d} is merged locally, its rows go to their owners,
//                             the owners apply the fused update to their shards; nothing is broadcast.
// Both call the C ABI's row-sharded engine (b200emb_shard_*, include/b200emb.h), one engine per device, held
// in a ShardedTableContext that the pass creates once per parameter (it owns no tensors: shards and optimizer
// state stay the variables of the local scopes).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "b200emb.h"
#include "paddle/fluid/framework/details/op_handle_base.h"
#include "paddle/fluid/framework/lod_tensor.h"
#include "paddle/fluid/framework/scope.h"
#include "paddle/fluid/framework/selected_rows.h"

namespace paddle {
namespace framework {
namespace details {

// One row-sharded parameter: the engines of all devices, connected to each other.
class ShardedTableContext {
 public:
  // shards[d] = device d's rows [ceil(height/G), D] (row i = global row i*G + d); moments likewise for Adam.
  // places may name the same GPU several times (fake devices, as reduce_op_handle_test.cc does with CPU
  // places): the handles then issue every phase for all ranks before the next one.
  ShardedTableContext(const std::vector<platform::Place>& places, int64_t height, int64_t d, int optimizer,
                      const std::vector<float*>& shards, const std::vector<float*>& moment1,
                      const std::vector<float*>& moment2, int64_t max_ids_per_step, int64_t max_recv_rows);
  ~ShardedTableContext();
  int world() const { return static_cast<int>(engines_.size()); }
  b200emb_shard_t* engine(int d) const { return engines_[d]; }
  const platform::Place& place(int d) const { return places_[d]; }
  int64_t height() const { return height_; }
  int64_t width() const { return d_; }
  bool shares_a_gpu() const { return shares_gpu_; }
  int NextSlot() { return slot_ = (slot_ + 1) % 4; }
  int CurrentSlot() const { return slot_; }

 private:
  std::vector<platform::Place> places_;
  std::vector<b200emb_shard_t*> engines_;
  int64_t height_, d_;
  bool shares_gpu_ = false;
  int slot_ = -1;
};

struct ShardedLookupOpHandle : public OpHandleBase {
  // inputs: one VarHandle per device naming Ids; outputs: one per device naming Out
  ShardedLookupOpHandle(const std::vector<Scope*>& local_scopes, ShardedTableContext* table)
      : local_scopes_(local_scopes), table_(table) {}
  std::string Name() const override { return "sharded_lookup"; }
  bool IsMultiDeviceTransfer() override { return true; }

 protected:
  void RunImpl() override;
  std::vector<Scope*> local_scopes_;
  ShardedTableContext* table_;
};

struct SparseAllToAllOpHandle : public OpHandleBase {
  // inputs: per device the sparse gradient W@GRAD (SelectedRows whose rows are the ids of the preceding
  // ShardedLookupOpHandle, as lookup_table_grad produces them) and then per device LearningRate (+ Beta1Pow,
  // Beta2Pow for Adam); no outputs besides the shards updated in place
  SparseAllToAllOpHandle(const std::vector<Scope*>& local_scopes, ShardedTableContext* table, float beta1 = 0.9f,
                         float beta2 = 0.999f, float epsilon = 1e-8f)
      : local_scopes_(local_scopes), table_(table), beta1_(beta1), beta2_(beta2), epsilon_(epsilon) {}
  std::string Name() const override { return "sparse_all_to_all"; }
  bool IsMultiDeviceTransfer() override { return true; }

 protected:
  void RunImpl() override;
  std::vector<Scope*> local_scopes_;
  ShardedTableContext* table_;
  float beta1_, beta2_, epsilon_;
};

}  // namespace details
}  // namespace framework
}  // namespace paddle
