// lookup_table_op.cu — replaces paddle/fluid/operators/lookup_table_op.cu of the reference.
// Same op names, slots and attrs (lookup_table_op.cc:59-121); kernels are the B200 library's.
#include "b200_op_common.h"
#include "paddle/fluid/framework/lod_tensor.h"

namespace paddle {
namespace operators {

using framework::LoDTensor;
using framework::SelectedRows;

template <typename T>
class LookupTableB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& context) const override {
    auto* table_t = context.Input<LoDTensor>("W");
    auto* ids_t = context.Input<LoDTensor>("Ids");
    auto* output_t = context.Output<LoDTensor>("Out");
    const int64_t padding_idx = context.Attr<int64_t>("padding_idx");
    const int64_t V = table_t->dims()[0], D = table_t->dims()[1], K = ids_t->numel();
    // Out takes Ids' leading dims with the last one replaced by D (lookup_table_op.cc:38-45)
    auto out_dims = framework::vectorize(ids_t->dims());
    out_dims.back() = D;
    output_t->Resize(framework::make_ddim(out_dims));
    output_t->set_lod(ids_t->lod());                      // LoD forwarded (lookup_table_op.cc:46-49)
    T* out = output_t->mutable_data<T>(context.GetPlace());
    if (K == 0) return;
    auto stream = context.cuda_device_context().stream();
    b200::StatusWord status(context.GetPlace(), stream);
    B200EMB_CHECK(b200::Api<T>::lookup(table_t->data<T>(), V, D, ids_t->data<int64_t>(), K, padding_idx, out,
                                       status.get(), stream));
    // fused mode: the sparse gradient of this lookup will be sorted by these very ids — start that now, on the
    // side stream, so the optimizer op finds the plan ready (b200_side_table.h)
    if (b200::SideTable::Instance().fused() && context.Attr<bool>("is_sparse"))
      b200::SideTable::Instance().StartPlan(boost::get<platform::CUDAPlace>(context.GetPlace()).device,
                                            context.GetPlace(), ids_t->data<int64_t>(), K, V, D, sizeof(T), stream);
    status.Enforce("lookup_table");
  }
};

template <typename T>
class LookupTableGradB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& context) const override {
    const bool is_sparse = context.Attr<bool>("is_sparse");
    auto* table = context.Input<LoDTensor>("W");
    auto* ids = context.Input<LoDTensor>("Ids");
    auto* d_output = context.Input<LoDTensor>(framework::GradVarName("Out"));
    const int64_t V = table->dims()[0], D = table->dims()[1], K = ids->numel();
    auto stream = context.cuda_device_context().stream();
    if (is_sparse) {
      // rows = ids verbatim (duplicates kept), value = dOut flattened to [K, D]: lookup_table_op.h:106-134.
      // rows are written on the device only; no host round trip (cf. lookup_table_op.cu:125-134).
      auto* d_table = context.Output<SelectedRows>(framework::GradVarName("W"));
      auto* rows = d_table->mutable_rows();
      rows->resize(static_cast<size_t>(K));
      if (b200::SideTable::Instance().fused()) {
        // zero copy: the gradient IS (Ids, dOut); consumers read both through the side table, rows() keeps
        // its size only (b200_side_table.h).  The value tensor shares dOut's memory.
        d_table->mutable_value()->ShareDataWith(*d_output);
        d_table->mutable_value()->Resize({K, D});
        d_table->set_height(V);
        b200::SparseGradRecord rec;
        rec.ids = ids->data<int64_t>();
        rec.value = d_output->data<T>();
        rec.n = K; rec.d = D; rec.height = V;
        b200::SideTable::Instance().Bind(boost::get<platform::CUDAPlace>(context.GetPlace()).device, d_table, rec);
        return;
      }
      auto* d_table_value = d_table->mutable_value();
      d_table_value->Resize({K, D});
      T* value = d_table_value->mutable_data<T>(context.GetPlace());
      d_table->set_height(V);
      PADDLE_ENFORCE_EQ(d_table_value->dims(),
                        framework::flatten_to_2d(d_output->dims(), d_output->dims().size() - 1));
      if (K == 0) return;
      int64_t* rows_dev = rows->CUDAMutableData(context.GetPlace());
      B200EMB_CHECK(b200::Api<T>::grad_sparse(ids->data<int64_t>(), d_output->data<T>(), K, D, rows_dev, value, stream));
    } else {
      // dense gradient: zero [V, D] then add every row (lookup_table_op.h:135-157), without atomics
      auto* d_table = context.Output<LoDTensor>(framework::GradVarName("W"));
      d_table->Resize(table->dims());
      T* dt = d_table->mutable_data<T>(context.GetPlace());
      PADDLE_ENFORCE(cudaMemsetAsync(dt, 0, sizeof(T) * V * D, stream) == cudaSuccess);
      if (K == 0) return;
      framework::Tensor ws;
      size_t ws_bytes = 0;
      void* wsp = b200::Workspace(&ws, K, D, sizeof(T), context.GetPlace(), &ws_bytes);
      B200EMB_CHECK(b200::Api<T>::add_to_tensor(ids->data<int64_t>(), d_output->data<T>(), K, D, V, dt, wsp,
                                                ws_bytes, b200::Flags(), nullptr, stream));
      // `ws` dies here like any Fluid temporary: the allocator's reuse is ordered by the op stream
    }
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(lookup_table, ops::LookupTableB200Kernel<float>, ops::LookupTableB200Kernel<double>);
REGISTER_OP_CUDA_KERNEL(lookup_table_grad, ops::LookupTableGradB200Kernel<float>,
                        ops::LookupTableGradB200Kernel<double>);
