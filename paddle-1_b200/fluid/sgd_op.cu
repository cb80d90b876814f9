// sgd_op.cu — replaces paddle/fluid/operators/sgd_op.cu of the reference (same op, slots:
// Param, Grad, LearningRate -> ParamOut; sgd_op.cc:24-79).  The SelectedRows-gradient branch is
// the fused duplicate-merge + row update of libb200emb; the dense branch is the plain
// p_out = p - lr * g kernel the reference also has (sgd_op.cu:24-34), kept so the op stays complete.
#include "b200_op_common.h"

namespace paddle {
namespace operators {

namespace {
template <typename T>
__global__ void DenseSGD(const T* g, const T* p, const T* lr, int64_t n, T* p_out) {
  const T rate = lr[0];
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    p_out[i] = p[i] - rate * g[i];
}
}  // namespace

template <typename T>
class SGDOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto* param = ctx.Input<framework::Tensor>("Param");
    auto* param_out = ctx.Output<framework::Tensor>("ParamOut");
    auto* learning_rate = ctx.Input<framework::Tensor>("LearningRate");
    auto* grad_var = ctx.InputVar("Grad");
    auto stream = ctx.cuda_device_context().stream();
    if (grad_var->IsType<framework::LoDTensor>()) {
      param_out->Resize(param->dims());
      T* out = param_out->mutable_data<T>(ctx.GetPlace());
      auto* grad = ctx.Input<framework::Tensor>("Grad");
      const int64_t n = param->numel();
      if (n == 0) return;
      const int grid = static_cast<int>(std::min<int64_t>((n + 511) / 512, 148 * 8));
      DenseSGD<T><<<grid, 512, 0, stream>>>(grad->data<T>(), param->data<T>(), learning_rate->data<T>(), n, out);
    } else if (grad_var->IsType<framework::SelectedRows>()) {
      PADDLE_ENFORCE_EQ(param, param_out);                       // in place, sgd_op.cu:84
      auto* grad = ctx.Input<framework::SelectedRows>("Grad");
      const int device = boost::get<platform::CUDAPlace>(ctx.GetPlace()).device;
      const b200::SparseGradView g = b200::ViewSparseGrad<T>(device, ctx.GetPlace(), grad);
      if (g.n == 0) return;                                      // sgd_op.h:56-58 (the CUDA file forgot it)
      PADDLE_ENFORCE_EQ(g.height, param_out->dims()[0]);         // sgd_op.h:60-62
      PADDLE_ENFORCE_EQ(g.d, param_out->numel() / g.height);
      b200::UpdateWorkspace w;
      b200::PrepareWorkspace(&w, g, sizeof(T), ctx.GetPlace(), stream);
      b200::StatusWord status(ctx.GetPlace(), stream);
      B200EMB_CHECK(b200::Api<T>::sgd(param_out->data<T>(), g.height, g.d, g.rows, static_cast<const T*>(g.value), g.n,
                                      learning_rate->data<T>(), w.ptr, w.bytes, w.flags, status.get(), stream));
      b200::SideTable::Instance().Release(device, g.plan);
      status.Enforce("sgd");
    } else {
      PADDLE_THROW("Unsupported Variable Type of Grad");
    }
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(sgd, ops::SGDOpB200Kernel<float>, ops::SGDOpB200Kernel<double>);
