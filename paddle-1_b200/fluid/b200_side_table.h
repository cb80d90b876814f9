// b200_side_table.h — "fused mode" of the OpKernel wrappers (SURVEY §7 hard part 2).
//
// `framework::SelectedRows::rows_` is a host-first container: honouring it literally costs the sparse gradient an
// H2D copy and a device sync per producer and per consumer (the reference pays exactly that:
// operators/lookup_table_op.cu:114-147, framework/mixed_vector.h:308-345).  In fused mode the wrappers keep
// the gradient on the device instead:
//   * `lookup_table` (forward, is_sparse) starts the sort plan of its Ids on a per-device side stream — the ids
//     are all the plan needs — into a pooled workspace, keyed by the Ids' device pointer;
//   * `lookup_table_grad` does not copy anything: it records {ids pointer, dOut pointer, N, D, height} (and the
//     plan, if one was started for these ids) under the output SelectedRows variable;
//   * `sum` / `sgd` / `adam` / `adagrad` look their SelectedRows input up here first, read ids and values from
//     the recorded device pointers (rows() is never touched) and run with B200EMB_PLAN_READY when a plan exists.
// The mode is a process flag (FLAGS_b200emb_fused in a Paddle tree; SetFused() / env B200EMB_FUSED here), not
// auto-detected: a kernel sees only its own op.  It is valid when every consumer of the sparse gradient is one
// of these wrappers — true for lookup_table_grad -> [sum] -> sgd|adam|adagrad, false e.g. with extract_rows.
//
// State is process-global, per device, mutex-protected: a fresh kernel object is constructed per call
// (framework/op_registry.h:101-103) and ParallelExecutor runs ops from a thread pool
// (framework/details/threaded_ssa_graph_executor.cc:211-228).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "b200emb.h"
#include "paddle/fluid/framework/selected_rows.h"
#include "paddle/fluid/framework/tensor.h"

namespace paddle {
namespace operators {
namespace b200 {

struct PlanSlot {              // a pooled workspace holding (or about to hold) one sort plan
  framework::Tensor ws;
  size_t bytes = 0;
  cudaEvent_t done = nullptr;  // recorded on the side stream after b200emb_plan_rows
  int64_t n = 0, d = 0, height = 0;
  size_t elem = 0;
};

struct SparseGradRecord {
  const int64_t* ids = nullptr;   // device: the gradient's rows (= the lookup's Ids, duplicates kept)
  const void* value = nullptr;    // device: the gradient's value block [n, d]
  int64_t n = 0, d = 0, height = 0;
  std::shared_ptr<PlanSlot> plan; // sort plan of `ids`, or null
  std::shared_ptr<framework::Tensor> keep;   // device memory `ids` lives in when no variable owns it (sum's output)
};

class SideTable {
 public:
  static SideTable& Instance() {
    static SideTable t;
    return t;
  }
  bool fused() const { return fused_.load(std::memory_order_relaxed); }
  void SetFused(bool on) { fused_.store(on, std::memory_order_relaxed); }

  // forward op: build the sort plan of `ids` beside the main stream.  The side stream first waits for
  // everything queued on `main` so far (the ids may have been produced there).
  void StartPlan(int device, const platform::Place& place, const int64_t* ids, int64_t n, int64_t height,
                 int64_t d, size_t elem, cudaStream_t main) {
    if (n <= 0) return;
    Dev& dv = dev(device);
    std::lock_guard<std::mutex> lock(dv.mu);
    if (!dv.side) {
      cudaSetDevice(device);
      cudaStreamCreateWithFlags(&dv.side, cudaStreamNonBlocking);
      cudaEventCreateWithFlags(&dv.fork, cudaEventDisableTiming);
    }
    std::shared_ptr<PlanSlot> slot;
    if (!dv.pool.empty()) {
      slot = dv.pool.back();
      dv.pool.pop_back();
    } else {
      slot = std::make_shared<PlanSlot>();
      cudaEventCreateWithFlags(&slot->done, cudaEventDisableTiming);
    }
    const size_t need = b200emb_workspace_bytes(n, d, static_cast<int64_t>(elem));
    if (slot->bytes < need) {   // grows to the largest request seen, then stays (a persistent temporary)
      slot->ws.mutable_data<uint8_t>(framework::make_ddim({static_cast<int64_t>(need)}), place);
      slot->bytes = need;
    }
    slot->n = n; slot->d = d; slot->height = height; slot->elem = elem;
    cudaEventRecord(dv.fork, main);
    cudaStreamWaitEvent(dv.side, dv.fork, 0);
    if (b200emb_plan_rows(ids, n, height, slot->ws.data<uint8_t>(), slot->bytes, nullptr, dv.side) != B200EMB_OK) {
      dv.pool.push_back(slot);
      return;
    }
    cudaEventRecord(slot->done, dv.side);
    dv.plans[ids] = slot;
  }

  // grad op: remember where the sparse gradient of `var` lives; adopts the plan started for these ids
  void Bind(int device, const framework::SelectedRows* var, SparseGradRecord rec) {
    Dev& dv = dev(device);
    std::lock_guard<std::mutex> lock(dv.mu);
    auto it = dv.plans.find(rec.ids);
    if (it != dv.plans.end()) {
      if (it->second->n == rec.n && it->second->height == rec.height) rec.plan = it->second;
      else dv.pool.push_back(it->second);
      dv.plans.erase(it);
    }
    auto old = dv.records.find(var);
    if (old != dv.records.end() && old->second.plan) dv.pool.push_back(old->second.plan);
    dv.records[var] = rec;
  }

  // consumer: take the record of `var` (false: not a fused gradient — use rows()/value() as usual)
  bool Take(int device, const framework::SelectedRows* var, SparseGradRecord* out) {
    Dev& dv = dev(device);
    std::lock_guard<std::mutex> lock(dv.mu);
    auto it = dv.records.find(var);
    if (it == dv.records.end()) return false;
    *out = it->second;
    dv.records.erase(it);
    return true;
  }

  // consumer, after it has queued its kernels on the main stream: the slot may host the next plan (whose
  // kernels wait for the main stream as it is at that later StartPlan)
  void Release(int device, const std::shared_ptr<PlanSlot>& slot) {
    if (!slot) return;
    Dev& dv = dev(device);
    std::lock_guard<std::mutex> lock(dv.mu);
    dv.pool.push_back(slot);
  }

 private:
  struct Dev {
    std::mutex mu;
    cudaStream_t side = nullptr;
    cudaEvent_t fork = nullptr;
    std::unordered_map<const void*, std::shared_ptr<PlanSlot>> plans;            // by Ids device pointer
    std::unordered_map<const framework::SelectedRows*, SparseGradRecord> records;  // by gradient variable
    std::vector<std::shared_ptr<PlanSlot>> pool;
  };
  SideTable() {
    const char* e = std::getenv("B200EMB_FUSED");
    fused_.store(e && e[0] && e[0] != '0');
  }
  Dev& dev(int device) {
    std::lock_guard<std::mutex> lock(mu_);
    auto& p = devs_[device];
    if (!p) p.reset(new Dev());
    return *p;
  }
  std::mutex mu_;
  std::map<int, std::unique_ptr<Dev>> devs_;
  std::atomic<bool> fused_{false};
};

// What a consumer op needs of its SelectedRows gradient, from the side table when the gradient is a fused one,
// from the variable itself otherwise.
struct SparseGradView {
  const int64_t* rows = nullptr;
  const void* value = nullptr;
  int64_t n = 0, d = 0, height = 0;
  std::shared_ptr<PlanSlot> plan;   // non-null: pass plan->ws with B200EMB_PLAN_READY after waiting for plan->done
  std::shared_ptr<framework::Tensor> keep;   // keeps recorded row ids alive while the consumer's kernels are queued
  bool fused = false;
};

template <typename T>
inline SparseGradView ViewSparseGrad(int device, const platform::Place& place, const framework::SelectedRows* grad) {
  SparseGradView v;
  SparseGradRecord rec;
  if (SideTable::Instance().fused() && SideTable::Instance().Take(device, grad, &rec)) {
    v.rows = rec.ids; v.value = rec.value; v.n = rec.n; v.d = rec.d; v.height = rec.height;
    v.plan = rec.plan;
    v.keep = rec.keep;
    v.fused = true;
    return v;
  }
  v.n = static_cast<int64_t>(grad->rows().size());
  v.height = grad->height();
  if (v.n > 0) {
    v.d = grad->value().numel() / v.n;
    v.rows = grad->rows().CUDAData(place);
    v.value = grad->value().data<T>();
  }
  return v;
}

}  // namespace b200
}  // namespace operators
}  // namespace paddle
