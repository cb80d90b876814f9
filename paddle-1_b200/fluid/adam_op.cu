// adam_op.cu — replaces paddle/fluid/operators/adam_op.cu of the reference (op `adam`, slots and
// attrs of adam_op.cc:24-120).  SelectedRows gradient: MergeAdd fused with the lazy row update
// (adam_op.h:282-321) in libb200emb.  Dense gradient: the element-wise update with the same
// operation order as AdamFunctor (adam_op.h:63-90).
#include "b200_op_common.h"
#include "paddle/fluid/operators/detail/safe_ref.h"

namespace paddle {
namespace operators {

namespace {
template <typename T>
__global__ void DenseAdam(T beta1, T beta2, T epsilon, const T* beta1_pow_p, const T* beta2_pow_p, const T* mom1,
                          T* mom1_out, const T* mom2, T* mom2_out, const T* lr_p, const T* grad, const T* param,
                          T* param_out, int64_t n) {
  const T beta1_pow = *beta1_pow_p, beta2_pow = *beta2_pow_p;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    T g = grad[i], m1 = mom1[i], m2 = mom2[i], lr = *lr_p, p = param[i];
    lr *= sqrt(1 - beta2_pow) / (1 - beta1_pow);
    m1 = beta1 * m1 + (1 - beta1) * g;
    m2 = beta2 * m2 + (1 - beta2) * g * g;
    p -= lr * (m1 / (sqrt(m2) + epsilon));
    mom1_out[i] = m1;
    mom2_out[i] = m2;
    param_out[i] = p;
  }
}
}  // namespace

template <typename T>
class AdamOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    using paddle::framework::LoDTensor;
    using paddle::operators::detail::Ref;
    const float beta1 = ctx.Attr<float>("beta1"), beta2 = ctx.Attr<float>("beta2"), epsilon = ctx.Attr<float>("epsilon");
    auto& param = Ref(ctx.Input<LoDTensor>("Param"), "Must set Param");
    auto* grad_var = ctx.InputVar("Grad");
    auto& mom1 = Ref(ctx.Input<LoDTensor>("Moment1"), "Must set Moment1");
    auto& mom2 = Ref(ctx.Input<LoDTensor>("Moment2"), "Must set Moment2");
    auto& lr = Ref(ctx.Input<LoDTensor>("LearningRate"), "Must set LearningRate");
    auto& beta1_pow = Ref(ctx.Input<LoDTensor>("Beta1Pow"), "Must set Beta1Pow");
    auto& beta2_pow = Ref(ctx.Input<LoDTensor>("Beta2Pow"), "Must set Beta2Pow");
    auto& param_out = Ref(ctx.Output<LoDTensor>("ParamOut"), "Must set ParamOut");
    auto& mom1_out = Ref(ctx.Output<LoDTensor>("Moment1Out"), "Must set Moment1Out");
    auto& mom2_out = Ref(ctx.Output<LoDTensor>("Moment2Out"), "Must set Moment1Out");
    auto stream = ctx.cuda_device_context().stream();
    param_out.Resize(param.dims());
    mom1_out.Resize(mom1.dims());
    mom2_out.Resize(mom2.dims());
    T* p_out = param_out.template mutable_data<T>(ctx.GetPlace());
    T* m1_out = mom1_out.template mutable_data<T>(ctx.GetPlace());
    T* m2_out = mom2_out.template mutable_data<T>(ctx.GetPlace());
    if (grad_var->IsType<framework::LoDTensor>()) {
      auto& grad = Ref(ctx.Input<LoDTensor>("Grad"), "Must set Grad");
      const int64_t n = param.numel();
      if (n == 0) return;
      const int grid = static_cast<int>(std::min<int64_t>((n + 255) / 256, 148 * 8));
      DenseAdam<T><<<grid, 256, 0, stream>>>(static_cast<T>(beta1), static_cast<T>(beta2), static_cast<T>(epsilon),
                                             beta1_pow.template data<T>(), beta2_pow.template data<T>(),
                                             mom1.template data<T>(), m1_out, mom2.template data<T>(), m2_out,
                                             lr.template data<T>(), grad.template data<T>(), param.template data<T>(),
                                             p_out, n);
    } else if (grad_var->IsType<framework::SelectedRows>()) {
      auto& grad = Ref(ctx.Input<framework::SelectedRows>("Grad"), "Must set Grad");
      const int device = boost::get<platform::CUDAPlace>(ctx.GetPlace()).device;
      const b200::SparseGradView g = b200::ViewSparseGrad<T>(device, ctx.GetPlace(), &grad);
      if (g.n == 0) return;                                                    // adam_op.h:285-288
      // the sparse update is in place: only touched rows change (lazy Adam), so outputs must alias
      // their inputs, as python/paddle/fluid/optimizer.py:562-566 wires them
      PADDLE_ENFORCE_EQ(p_out, param.template data<T>(), "sparse adam needs ParamOut == Param");
      PADDLE_ENFORCE_EQ(m1_out, mom1.template data<T>(), "sparse adam needs Moment1Out == Moment1");
      PADDLE_ENFORCE_EQ(m2_out, mom2.template data<T>(), "sparse adam needs Moment2Out == Moment2");
      PADDLE_ENFORCE_EQ(g.height, param.dims()[0]);
      b200::UpdateWorkspace w;
      b200::PrepareWorkspace(&w, g, sizeof(T), ctx.GetPlace(), stream);
      b200::StatusWord status(ctx.GetPlace(), stream);
      B200EMB_CHECK(b200::Api<T>::adam(p_out, m1_out, m2_out, g.height, g.d, g.rows, static_cast<const T*>(g.value), g.n,
                                       lr.template data<T>(), beta1_pow.template data<T>(),
                                       beta2_pow.template data<T>(), beta1, beta2, epsilon, w.ptr, w.bytes,
                                       w.flags, status.get(), stream));
      b200::SideTable::Instance().Release(device, g.plan);
      status.Enforce("adam");
    } else {
      PADDLE_THROW("Variable type not supported by adam_op");
    }
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(adam, ops::AdamOpB200Kernel<float>, ops::AdamOpB200Kernel<double>);
