// runner.cu — a minimal stand-in for the part of Fluid that sits between Python and an op kernel:
// a Scope of Variables, `Operator.run(scope, place)` (cf. pybind.cc:469-490, operator.cc:657-714:
// pick the kernel by (op type, dtype of the key input, place) from AllOpKernels() and call it),
// exported as a small C API for ctypes.  It exists so the reference's Python op tests can be mirrored
// one to one against the wrappers in this directory; it is not an executor.
#include <cstring>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "paddle/fluid/framework/lod_tensor.h"
#include "paddle/fluid/framework/op_registry.h"
#include "paddle/fluid/framework/selected_rows.h"
#include "b200_side_table.h"
#include "sparse_all_to_all_op_handle.h"

namespace fw = paddle::framework;
namespace pf = paddle::platform;

namespace {
thread_local std::string g_err;

struct Session {
  fw::Scope scope;
  std::map<int, std::unique_ptr<pf::CUDADeviceContext>> ctxs;
  pf::CUDADeviceContext& ctx(int dev) {
    auto it = ctxs.find(dev);
    if (it == ctxs.end()) it = ctxs.emplace(dev, new pf::CUDADeviceContext(pf::CUDAPlace(dev))).first;
    return *it->second;
  }
};

// dtype codes: 0 f32, 1 f64, 2 i64, 3 i32
size_t ElemSize(int code) { return code == 0 ? 4 : code == 1 ? 8 : code == 2 ? 8 : 4; }

fw::Tensor* TensorOf(fw::Variable* var, bool create) {
  if (var->IsType<fw::SelectedRows>()) return var->GetMutable<fw::SelectedRows>()->mutable_value();
  if (var->IsType<fw::LoDTensor>() || create) return var->GetMutable<fw::LoDTensor>();
  return nullptr;
}

void* Alloc(fw::Tensor* t, int code, const pf::Place& place) {
  switch (code) {
    case 0: return t->mutable_data<float>(place);
    case 1: return t->mutable_data<double>(place);
    case 2: return t->mutable_data<int64_t>(place);
    default: return t->mutable_data<int>(place);
  }
}

int CodeOf(const fw::Tensor& t) {
  if (t.type() == std::type_index(typeid(float))) return 0;
  if (t.type() == std::type_index(typeid(double))) return 1;
  if (t.type() == std::type_index(typeid(int64_t))) return 2;
  return 3;
}

// "Slot=a,b;Other=c"
fw::VariableNameMap ParseSlots(const char* spec) {
  fw::VariableNameMap m;
  std::stringstream ss(spec ? spec : "");
  std::string item;
  while (std::getline(ss, item, ';')) {
    if (item.empty()) continue;
    const size_t eq = item.find('=');
    std::string slot = item.substr(0, eq), rest = item.substr(eq + 1);
    auto& v = m[slot];
    std::stringstream rs(rest);
    std::string name;
    while (std::getline(rs, name, ',')) if (!name.empty()) v.push_back(name);
  }
  return m;
}
// "name:kind:value;..."  kinds: b(ool) i(nt) l(ong) f(loat) s(tring) v(ector<int>, '|'-separated)
fw::AttributeMap ParseAttrs(const char* spec) {
  fw::AttributeMap m;
  std::stringstream ss(spec ? spec : "");
  std::string item;
  while (std::getline(ss, item, ';')) {
    if (item.empty()) continue;
    const size_t a = item.find(':'), b = item.find(':', a + 1);
    const std::string name = item.substr(0, a), kind = item.substr(a + 1, b - a - 1), val = item.substr(b + 1);
    if (kind == "b") m[name] = (val == "1");
    else if (kind == "i") m[name] = std::stoi(val);
    else if (kind == "l") m[name] = static_cast<int64_t>(std::stoll(val));
    else if (kind == "f") m[name] = std::stof(val);
    else if (kind == "v") {
      std::vector<int> v;
      std::stringstream vs(val);
      std::string x;
      while (std::getline(vs, x, '|')) if (!x.empty()) v.push_back(std::stoi(x));
      m[name] = v;
    }
    else m[name] = val;
  }
  return m;
}

// which input decides the kernel's data type (GetExpectedKernelType of each op)
const char* KeySlot(const std::string& type) {
  if (type == "lookup_table" || type == "lookup_table_grad") return "W";   // lookup_table_op.cc:53-57
  if (type == "sgd" || type == "adam" || type == "adagrad") return "Param";  // sgd_op.cc:44-48, adam_op.cc:76-81
  if (type == "split_ids") return "Ids";                                   // split_ids_op.cc:57-63
  return "X";                                                              // sum_op.cc:67-127, merge_ids_op.cc:101-107
}
}  // namespace

#define FL_TRY try {
#define FL_CATCH(ret) } catch (const std::exception& e) { g_err = e.what(); return ret; }

extern "C" {

const char* fl_last_error() { return g_err.c_str(); }
void* fl_session_new() { return new Session(); }
void fl_session_free(void* s) { delete static_cast<Session*>(s); }

// creates (or retypes) variable `name` as LoDTensor / SelectedRows
int fl_var_tensor(void* s, const char* name) {
  FL_TRY static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::LoDTensor>(); return 0; FL_CATCH(-1)
}
int fl_var_selected_rows(void* s, const char* name) {
  FL_TRY static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>(); return 0; FL_CATCH(-1)
}
int fl_var_kind(void* s, const char* name) {   // 0 missing/empty, 1 LoDTensor, 2 SelectedRows
  auto* v = static_cast<Session*>(s)->scope.FindVar(name);
  if (!v || !v->IsInitialized()) return 0;
  return v->IsType<fw::SelectedRows>() ? 2 : 1;
}

// tensor.set(np_array, place): device < 0 means CPUPlace
int fl_tensor_set(void* s, const char* name, int code, const int64_t* dims, int ndim, const void* host, int device) {
  FL_TRY
  auto* sess = static_cast<Session*>(s);
  fw::Tensor* t = TensorOf(sess->scope.Var(name), true);
  t->Resize(fw::make_ddim(std::vector<int64_t>(dims, dims + ndim)));
  const pf::Place place = device < 0 ? pf::Place(pf::CPUPlace()) : pf::Place(pf::CUDAPlace(device));
  void* dst = Alloc(t, code, place);
  const size_t bytes = static_cast<size_t>(t->numel()) * ElemSize(code);
  if (bytes) {
    if (device < 0) std::memcpy(dst, host, bytes);
    else PADDLE_ENFORCE(cudaMemcpy(dst, host, bytes, cudaMemcpyHostToDevice) == cudaSuccess);
  }
  return 0;
  FL_CATCH(-1)
}
// returns ndim (dims written), dtype code in *code; -1 on error, -2 when the tensor holds no memory
int fl_tensor_info(void* s, const char* name, int64_t* dims, int* code) {
  FL_TRY
  auto* v = static_cast<Session*>(s)->scope.FindVar(name);
  PADDLE_ENFORCE(v != nullptr, "no variable %s", name);
  fw::Tensor* t = TensorOf(v, false);
  PADDLE_ENFORCE(t != nullptr, "variable %s holds no tensor", name);
  for (int i = 0; i < t->dims().size(); ++i) dims[i] = t->dims()[i];
  if (!t->IsInitialized()) return -2;
  *code = CodeOf(*t);
  return t->dims().size();
  FL_CATCH(-1)
}
int fl_tensor_get(void* s, const char* name, void* host) {
  FL_TRY
  auto* v = static_cast<Session*>(s)->scope.FindVar(name);
  PADDLE_ENFORCE(v != nullptr, "no variable %s", name);
  fw::Tensor* t = TensorOf(v, false);
  PADDLE_ENFORCE(t != nullptr && t->IsInitialized(), "variable %s holds no data", name);
  const size_t bytes = static_cast<size_t>(t->numel()) * ElemSize(CodeOf(*t));
  if (!bytes) return 0;
  const void* src = t->data<char>();
  if (t->place().is_gpu()) {
    PADDLE_ENFORCE(cudaDeviceSynchronize() == cudaSuccess, "device error: %s", cudaGetErrorString(cudaGetLastError()));
    PADDLE_ENFORCE(cudaMemcpy(host, src, bytes, cudaMemcpyDeviceToHost) == cudaSuccess);
  } else {
    std::memcpy(host, src, bytes);
  }
  return 0;
  FL_CATCH(-1)
}

int fl_sr_set_rows(void* s, const char* name, const int64_t* rows, int64_t n) {
  FL_TRY
  auto* sr = static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>();
  sr->set_rows(std::vector<int64_t>(rows, rows + n));
  return 0;
  FL_CATCH(-1)
}
int fl_sr_set_height(void* s, const char* name, int64_t h) {
  FL_TRY static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>()->set_height(h); return 0; FL_CATCH(-1)
}
int64_t fl_sr_height(void* s, const char* name) {
  FL_TRY return static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>()->height(); FL_CATCH(-1)
}
int64_t fl_sr_num_rows(void* s, const char* name) {
  FL_TRY
  return static_cast<int64_t>(static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>()->rows().size());
  FL_CATCH(-1)
}
int fl_sr_get_rows(void* s, const char* name, int64_t* out) {
  FL_TRY
  PADDLE_ENFORCE(cudaDeviceSynchronize() == cudaSuccess);
  const auto& rows = static_cast<Session*>(s)->scope.Var(name)->GetMutable<fw::SelectedRows>()->rows();
  size_t i = 0;
  for (auto it = rows.begin(); it != rows.end(); ++it) out[i++] = *it;   // host read: downloads if the device copy is newer
  return 0;
  FL_CATCH(-1)
}

// Operator(type, ...).run(scope, place)
int fl_op_run(void* s, const char* type, const char* inputs, const char* outputs, const char* attrs, int device) {
  FL_TRY
  auto* sess = static_cast<Session*>(s);
  fw::OperatorBase op(type, ParseSlots(inputs), ParseSlots(outputs), ParseAttrs(attrs));
  auto kit = fw::AllOpKernels().find(type);
  PADDLE_ENFORCE(kit != fw::AllOpKernels().end(), "There are no kernels which are registered in the %s operator.", type);
  // data type from the key input (first initialised one for multi-input slots)
  std::type_index dtype(typeid(void));
  for (auto& vn : op.Inputs(KeySlot(type))) {
    auto* v = sess->scope.FindVar(vn);
    if (!v || !v->IsInitialized()) continue;
    fw::Tensor* t = TensorOf(v, false);
    if (t && t->IsInitialized()) { dtype = t->type(); break; }
  }
  PADDLE_ENFORCE(dtype != std::type_index(typeid(void)), "Sum operator should have at least one tensor / no key input for %s", type);
  PADDLE_ENFORCE(device >= 0, "op %s has no CPU kernel in this library (no CPU fallback)", type);
  auto fit = kit->second.find(fw::OpKernelKey{dtype, true});
  PADDLE_ENFORCE(fit != kit->second.end(), "op %s does not have kernel for this data type on CUDAPlace", type);
  // every output variable exists before the kernel runs (Executor::CreateVariables)
  for (auto& kv : op.Outputs()) for (auto& vn : kv.second) sess->scope.Var(vn);
  PADDLE_ENFORCE(cudaSetDevice(device) == cudaSuccess);                       // operator.cc:145-146
  auto& ctx = sess->ctx(device);
  fit->second(fw::ExecutionContext(op, sess->scope, ctx));
  ctx.Wait();                                                                 // executor.cc:340
  cudaError_t e = cudaGetLastError();
  PADDLE_ENFORCE(e == cudaSuccess, "CUDA error after %s: %s", type, cudaGetErrorString(e));
  return 0;
  FL_CATCH(-1)
}

// FLAGS_b200emb_fused of a real tree: sparse gradients stay on the device between our ops (b200_side_table.h)
int fl_set_fused(int on) {
  paddle::operators::b200::SideTable::Instance().SetFused(on != 0);
  return 0;
}

int fl_num_registered_kernels(const char* type) {
  auto it = fw::AllOpKernels().find(type);
  return it == fw::AllOpKernels().end() ? 0 : static_cast<int>(it->second.size());
}

}  // extern "C"

// ---- a stand-in for the part of ParallelExecutor the sharded-table op handles need: N local scopes, one
// device context per GPU (ranks that share a GPU share its stream), the handles wired to per-device variables
namespace {
namespace fd = paddle::framework::details;
struct ParallelSession {
  std::vector<std::unique_ptr<Session>> locals;
  std::vector<pf::Place> places;
  std::map<int, std::unique_ptr<pf::CUDADeviceContext>> ctxs;
  std::unique_ptr<fd::ShardedTableContext> table;
  std::vector<std::unique_ptr<fd::VarHandle>> handles;   // kept alive for the op handles of the last run
  std::vector<fw::Scope*> scopes() {
    std::vector<fw::Scope*> v;
    for (auto& l : locals) v.push_back(&l->scope);
    return v;
  }
  void wire(fd::OpHandleBase* op) {
    for (size_t i = 0; i < places.size(); ++i)
      op->SetDeviceContext(places[i], ctxs[boost::get<pf::CUDAPlace>(places[i]).device].get());
  }
  fd::VarHandle* var(size_t scope_idx, const char* name) {
    handles.emplace_back(new fd::VarHandle(0, scope_idx, name, places[scope_idx]));
    return handles.back().get();
  }
};
}  // namespace

extern "C" {

void* fl_pe_new(int n, const int* devices) {
  auto* pe = new ParallelSession();
  for (int i = 0; i < n; ++i) {
    pe->locals.emplace_back(new Session());
    pe->places.push_back(pf::CUDAPlace(devices[i]));
    if (!pe->ctxs.count(devices[i])) pe->ctxs[devices[i]].reset(new pf::CUDADeviceContext(pf::CUDAPlace(devices[i])));
  }
  return pe;
}
void fl_pe_free(void* p) { delete static_cast<ParallelSession*>(p); }
void* fl_pe_scope(void* p, int i) { return static_cast<ParallelSession*>(p)->locals[i].get(); }

// the pass's set-up for one embedding parameter: `param` (and moments) in local scope d hold device d's shard
int fl_pe_shard_table(void* p, const char* param, const char* m1, const char* m2, int64_t height, int64_t d,
                      int optimizer, int64_t max_ids, int64_t max_recv) {
  FL_TRY
  auto* pe = static_cast<ParallelSession*>(p);
  std::vector<float*> shards, mom1, mom2;
  for (size_t i = 0; i < pe->locals.size(); ++i) {
    auto get = [&](const char* name) {
      auto* v = pe->locals[i]->scope.FindVar(name);
      PADDLE_ENFORCE_NOT_NULL(v, "variable %s not found in local scope %d", name, static_cast<int>(i));
      return v->GetMutable<fw::LoDTensor>()->data<float>();
    };
    shards.push_back(get(param));
    if (optimizer == B200EMB_SHARD_ADAM) { mom1.push_back(get(m1)); mom2.push_back(get(m2)); }
  }
  pe->table.reset(new fd::ShardedTableContext(pe->places, height, d, optimizer, shards, mom1, mom2, max_ids, max_recv));
  return 0;
  FL_CATCH(-1)
}

int fl_pe_sharded_lookup(void* p, const char* ids, const char* out) {
  FL_TRY
  auto* pe = static_cast<ParallelSession*>(p);
  PADDLE_ENFORCE_NOT_NULL(pe->table.get(), "fl_pe_shard_table first");
  pe->handles.clear();
  fd::ShardedLookupOpHandle op(pe->scopes(), pe->table.get());
  pe->wire(&op);
  for (size_t i = 0; i < pe->places.size(); ++i) op.AddInput(pe->var(i, ids));
  for (size_t i = 0; i < pe->places.size(); ++i) op.AddOutput(pe->var(i, out));
  op.Run(true);
  return 0;
  FL_CATCH(-1)
}

int fl_pe_sparse_all_to_all(void* p, const char* grad, const char* lr, const char* b1p, const char* b2p, float beta1,
                            float beta2, float epsilon) {
  FL_TRY
  auto* pe = static_cast<ParallelSession*>(p);
  PADDLE_ENFORCE_NOT_NULL(pe->table.get(), "fl_pe_shard_table first");
  pe->handles.clear();
  fd::SparseAllToAllOpHandle op(pe->scopes(), pe->table.get(), beta1, beta2, epsilon);
  pe->wire(&op);
  const char* names[4] = {grad, lr, b1p, b2p};
  const int groups = (b1p && b1p[0]) ? 4 : 2;
  for (int k = 0; k < groups; ++k)
    for (size_t i = 0; i < pe->places.size(); ++i) op.AddInput(pe->var(i, names[k]));
  op.Run(true);
  return 0;
  FL_CATCH(-1)
}

// ParallelExecutor's end-of-run sync (every device context) + the engines' status words
int fl_pe_sync(void* p) {
  FL_TRY
  auto* pe = static_cast<ParallelSession*>(p);
  for (auto& kv : pe->ctxs) {
    PADDLE_ENFORCE(cudaSetDevice(kv.first) == cudaSuccess);
    kv.second->Wait();
    cudaError_t e = cudaGetLastError();
    PADDLE_ENFORCE(e == cudaSuccess, "CUDA error: %s", cudaGetErrorString(e));
  }
  int bits = 0;
  if (pe->table)
    for (int d = 0; d < pe->table->world(); ++d) {
      int32_t st = 0;
      PADDLE_ENFORCE(cudaMemcpy(&st, b200emb_shard_status_ptr(pe->table->engine(d)), sizeof(st), cudaMemcpyDeviceToHost) == cudaSuccess);
      bits |= st;
    }
  return bits;
  FL_CATCH(-1)
}

}  // extern "C"
