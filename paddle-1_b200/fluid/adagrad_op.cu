// adagrad_op.cu — replaces paddle/fluid/operators/adagrad_op.cu of the reference (op `adagrad`:
// Param, Grad, Moment, LearningRate -> ParamOut, MomentOut; attr epsilon; adagrad_op.cc:24-100).
// SelectedRows gradient: MergeAdd + moment/param row update fused in libb200emb (CPU semantics
// adagrad_op.cc:108-141).  Dense gradient: element-wise, same formula as adagrad_op.h:49-60.
#include "b200_op_common.h"

namespace paddle {
namespace operators {

namespace {
template <typename T>
__global__ void DenseAdagrad(const T* g, const T* m, const T* p, const T* lr, T epsilon, int64_t n, T* m_out,
                             T* p_out) {
  const T rate = lr[0];
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const T mo = m[i] + g[i] * g[i];
    m_out[i] = mo;
    p_out[i] = p[i] - rate * g[i] / (sqrt(mo) + epsilon);
  }
}
template <typename T> struct AdagradApi;
template <> struct AdagradApi<float> { static constexpr auto fn = b200emb_sparse_adagrad_f32; };
template <> struct AdagradApi<double> { static constexpr auto fn = b200emb_sparse_adagrad_f64; };
}  // namespace

template <typename T>
class AdagradOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto* param_out = ctx.Output<framework::Tensor>("ParamOut");
    auto* moment_out = ctx.Output<framework::Tensor>("MomentOut");
    auto* param = ctx.Input<framework::Tensor>("Param");
    auto* moment = ctx.Input<framework::Tensor>("Moment");
    auto* lr = ctx.Input<framework::Tensor>("LearningRate");
    const float epsilon = ctx.Attr<float>("epsilon");
    auto* grad_var = ctx.InputVar("Grad");
    auto stream = ctx.cuda_device_context().stream();
    param_out->Resize(param->dims());
    moment_out->Resize(moment->dims());
    T* p_out = param_out->mutable_data<T>(ctx.GetPlace());
    T* m_out = moment_out->mutable_data<T>(ctx.GetPlace());
    if (grad_var->IsType<framework::LoDTensor>()) {
      auto* grad = ctx.Input<framework::Tensor>("Grad");
      const int64_t n = param->numel();
      if (n == 0) return;
      const int grid = static_cast<int>(std::min<int64_t>((n + 255) / 256, 148 * 8));
      DenseAdagrad<T><<<grid, 256, 0, stream>>>(grad->data<T>(), moment->data<T>(), param->data<T>(), lr->data<T>(),
                                                static_cast<T>(epsilon), n, m_out, p_out);
    } else if (grad_var->IsType<framework::SelectedRows>()) {
      auto* grad = ctx.Input<framework::SelectedRows>("Grad");
      PADDLE_ENFORCE_EQ(param, param_out);          // adagrad_op.h:63-66
      PADDLE_ENFORCE_EQ(moment, moment_out);
      const int device = boost::get<platform::CUDAPlace>(ctx.GetPlace()).device;
      const b200::SparseGradView g = b200::ViewSparseGrad<T>(device, ctx.GetPlace(), grad);
      if (g.n == 0) return;
      PADDLE_ENFORCE_EQ(g.height, param_out->dims()[0]);
      b200::UpdateWorkspace w;
      b200::PrepareWorkspace(&w, g, sizeof(T), ctx.GetPlace(), stream);
      b200::StatusWord status(ctx.GetPlace(), stream);
      B200EMB_CHECK(AdagradApi<T>::fn(p_out, m_out, g.height, g.d, g.rows, static_cast<const T*>(g.value), g.n,
                                      lr->data<T>(), epsilon, w.ptr, w.bytes, w.flags, status.get(), stream));
      b200::SideTable::Instance().Release(device, g.plan);
      status.Enforce("adagrad");
    } else {
      PADDLE_THROW("Unsupported Variable Type of Grad");
    }
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(adagrad, ops::AdagradOpB200Kernel<float>, ops::AdagradOpB200Kernel<double>);
