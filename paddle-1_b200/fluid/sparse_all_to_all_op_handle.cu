// sparse_all_to_all_op_handle.cu — see the header.  Host-side sequencing only: every kernel is the C ABI's.
#include "sparse_all_to_all_op_handle.h"

#include <cuda_runtime.h>

#include "b200_side_table.h"

namespace paddle {
namespace framework {
namespace details {

namespace {
int DeviceOf(const platform::Place& p) { return boost::get<platform::CUDAPlace>(p).device; }

cudaStream_t StreamOf(OpHandleBase* op, const platform::Place& p) {
  auto* ctx = dynamic_cast<const platform::CUDADeviceContext*>(op->DeviceContext(p));
  PADDLE_ENFORCE_NOT_NULL(ctx, "no CUDADeviceContext set for the place of a sharded-table op handle");
  return ctx->stream();
}
}  // namespace

ShardedTableContext::ShardedTableContext(const std::vector<platform::Place>& places, int64_t height, int64_t d,
                                         int optimizer, const std::vector<float*>& shards,
                                         const std::vector<float*>& moment1, const std::vector<float*>& moment2,
                                         int64_t max_ids_per_step, int64_t max_recv_rows)
    : places_(places), height_(height), d_(d) {
  const int w = static_cast<int>(places.size());
  PADDLE_ENFORCE_EQ(static_cast<int>(shards.size()), w);
  engines_.assign(w, nullptr);
  for (int a = 0; a < w; ++a)
    for (int b = a + 1; b < w; ++b) shares_gpu_ = shares_gpu_ || DeviceOf(places[a]) == DeviceOf(places[b]);
  std::vector<void*> blocks(w);
  for (int r = 0; r < w; ++r) {
    PADDLE_ENFORCE(cudaSetDevice(DeviceOf(places[r])) == cudaSuccess);
    for (int o = 0; o < w; ++o) {   // one process: plain peer pointers instead of CUDA IPC
      if (DeviceOf(places[o]) == DeviceOf(places[r])) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(DeviceOf(places[o]), 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else PADDLE_ENFORCE(e == cudaSuccess, "no peer access between the devices of a sharded table");
    }
    const int rc = b200emb_shard_create(w, r, height, d, optimizer, shards[r],
                                        moment1.empty() ? nullptr : moment1[r], moment2.empty() ? nullptr : moment2[r],
                                        max_ids_per_step, max_recv_rows, &engines_[r]);
    PADDLE_ENFORCE_EQ(rc, 0, "b200emb_shard_create failed (%d)", rc);
    blocks[r] = b200emb_shard_block(engines_[r]);
  }
  for (int r = 0; r < w; ++r) {
    PADDLE_ENFORCE(cudaSetDevice(DeviceOf(places[r])) == cudaSuccess);
    const int rc = b200emb_shard_connect_ptrs(engines_[r], blocks.data());
    PADDLE_ENFORCE_EQ(rc, 0, "b200emb_shard_connect_ptrs failed (%d)", rc);
  }
}

ShardedTableContext::~ShardedTableContext() {
  for (size_t r = 0; r < engines_.size(); ++r) {
    if (!engines_[r]) continue;
    cudaSetDevice(DeviceOf(places_[r]));
    b200emb_shard_destroy(engines_[r]);
  }
}

void ShardedLookupOpHandle::RunImpl() {
  const int w = table_->world();
  PADDLE_ENFORCE_EQ(static_cast<int>(inputs_.size()), w, "one Ids handle per device");
  PADDLE_ENFORCE_EQ(static_cast<int>(outputs_.size()), w, "one Out handle per device");
  const int slot = table_->NextSlot();
  std::vector<float*> out(w, nullptr);
  std::vector<cudaStream_t> stream(w);
  for (int d = 0; d < w; ++d) {
    auto* in = dynamic_cast<VarHandle*>(inputs_[d]);
    auto* ov = dynamic_cast<VarHandle*>(outputs_[d]);
    PADDLE_ENFORCE(in != nullptr && ov != nullptr);
    Scope* scope = local_scopes_[in->scope_idx_];
    auto* ids_var = scope->FindVar(in->name_);
    PADDLE_ENFORCE_NOT_NULL(ids_var, "Ids variable %s not found on device %d", in->name_.c_str(), d);
    auto& ids = ids_var->Get<LoDTensor>();
    auto* out_t = scope->Var(ov->name_)->GetMutable<LoDTensor>();
    auto dims = vectorize(ids.dims());               // Out takes Ids' leading dims with the last one replaced by D
    dims.back() = table_->width();                   // (lookup_table_op.cc:38-45)
    out_t->Resize(make_ddim(dims));
    out_t->set_lod(ids.lod());
    out[d] = out_t->mutable_data<float>(table_->place(d));
    stream[d] = StreamOf(this, table_->place(d));
    PADDLE_ENFORCE(cudaSetDevice(DeviceOf(table_->place(d))) == cudaSuccess);
    const int rc = b200emb_shard_prepare(table_->engine(d), slot, ids.data<int64_t>(), ids.numel(), stream[d]);
    PADDLE_ENFORCE_EQ(rc, 0, "b200emb_shard_prepare failed (%d)", rc);
  }
  // kernels of different ranks wait for each other: ranks that share a GPU get every phase issued for all of
  // them before the next phase; real devices run their whole sequence concurrently
  if (table_->shares_a_gpu()) {
    for (int ph = 0; ph < 4; ++ph)
      for (int d = 0; d < w; ++d) {
        PADDLE_ENFORCE(cudaSetDevice(DeviceOf(table_->place(d))) == cudaSuccess);
        PADDLE_ENFORCE_EQ(b200emb_shard_lookup_phases(table_->engine(d), slot, out[d], stream[d], ph, ph), 0);
      }
  } else {
    for (int d = 0; d < w; ++d) {
      PADDLE_ENFORCE(cudaSetDevice(DeviceOf(table_->place(d))) == cudaSuccess);
      PADDLE_ENFORCE_EQ(b200emb_shard_lookup(table_->engine(d), slot, out[d], stream[d]), 0);
    }
  }
}

void SparseAllToAllOpHandle::RunImpl() {
  const int w = table_->world();
  const bool adam = static_cast<int>(inputs_.size()) == 4 * w;
  PADDLE_ENFORCE(adam || static_cast<int>(inputs_.size()) == 2 * w,
                 "inputs: W@GRAD and LearningRate per device (+ Beta1Pow, Beta2Pow for adam)");
  const int slot = table_->CurrentSlot();
  PADDLE_ENFORCE_GE(slot, 0, "sparse_all_to_all runs after the sharded_lookup of the same step");
  std::vector<const float*> value(w), lr(w), b1p(w, nullptr), b2p(w, nullptr);
  std::vector<cudaStream_t> stream(w);
  for (int d = 0; d < w; ++d) {
    auto* gh = dynamic_cast<VarHandle*>(inputs_[d]);
    PADDLE_ENFORCE_NOT_NULL(gh);
    Scope* scope = local_scopes_[gh->scope_idx_];
    auto* gvar = scope->FindVar(gh->name_);
    PADDLE_ENFORCE(gvar != nullptr && gvar->IsType<SelectedRows>(), "sparse_all_to_all needs a SelectedRows gradient");
    auto& grad = gvar->Get<SelectedRows>();
    // lookup_table_grad takes the height from its W input, which on a sharded replica is the shard
    const int64_t shard_h = (table_->height() + w - 1) / w;
    PADDLE_ENFORCE(grad.height() == table_->height() || grad.height() == shard_h,
                   "gradient height %ld matches neither the table nor its shard", static_cast<long>(grad.height()));
    // the gradient of a sharded lookup: rows = the ids of that lookup (lookup_table_op.h:114-119), which the
    // engine already holds de-duplicated in the step's slot; only the value block is needed here
    operators::b200::SparseGradRecord rec;
    if (operators::b200::SideTable::Instance().fused() &&
        operators::b200::SideTable::Instance().Take(DeviceOf(table_->place(d)), &grad, &rec)) {
      value[d] = static_cast<const float*>(rec.value);
      operators::b200::SideTable::Instance().Release(DeviceOf(table_->place(d)), rec.plan);
    } else {
      value[d] = grad.rows().size() ? grad.value().data<float>() : nullptr;
    }
    auto get1 = [&](int k) {
      auto* h = dynamic_cast<VarHandle*>(inputs_[k * w + d]);
      PADDLE_ENFORCE_NOT_NULL(h);
      auto* v = local_scopes_[h->scope_idx_]->FindVar(h->name_);
      PADDLE_ENFORCE_NOT_NULL(v, "variable %s not found", h->name_.c_str());
      return v->Get<LoDTensor>().data<float>();
    };
    lr[d] = get1(1);
    if (adam) { b1p[d] = get1(2); b2p[d] = get1(3); }
    stream[d] = StreamOf(this, table_->place(d));
  }
  const int phases = table_->shares_a_gpu() ? 2 : 1;
  for (int ph = 0; ph < phases; ++ph)
    for (int d = 0; d < w; ++d) {
      PADDLE_ENFORCE(cudaSetDevice(DeviceOf(table_->place(d))) == cudaSuccess);
      const int first = phases == 2 ? ph : 0, last = phases == 2 ? ph : 1;
      const int rc = b200emb_shard_update_phases(table_->engine(d), slot, value[d], lr[d], b1p[d], b2p[d], beta1_,
                                                 beta2_, epsilon_, stream[d], first, last);
      PADDLE_ENFORCE_EQ(rc, 0, "b200emb_shard_update failed (%d)", rc);
    }
}

}  // namespace details
}  // namespace framework
}  // namespace paddle
