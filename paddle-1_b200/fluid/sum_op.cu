// sum_op.cu — replaces paddle/fluid/operators/sum_op.cu of the reference (op `sum`, X[] -> Out).
// SelectedRows output: concatenation of the non-empty inputs (sum_op.h:71-145 / SelectedRowsAddTo).
// LoDTensor output: dense inputs are added element-wise, SelectedRows inputs are accumulated with
// the duplicate-merging kernel (sum_op.h:40-70 / SelectedRowsAddToTensor).
#include <memory>
#include <type_traits>
#include <vector>
#include "b200_op_common.h"
#include "paddle/fluid/framework/lod_tensor.h"

namespace paddle {
namespace operators {

namespace {
template <typename T>
__global__ void AddInto(const T* x, int64_t n, T* out) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    out[i] = out[i] + x[i];
}
}  // namespace

template <typename T>
class SumB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& context) const override {
    using framework::LoDTensor;
    using framework::SelectedRows;
    auto in_vars = context.MultiInputVar("X");
    const int N = static_cast<int>(in_vars.size());
    auto* out_var = context.OutputVar("Out");
    const bool in_place = out_var == in_vars[0];
    auto stream = context.cuda_device_context().stream();
    const auto place = context.GetPlace();
    const int device = boost::get<platform::CUDAPlace>(place).device;

    if (out_var->IsType<LoDTensor>()) {
      auto* out = context.Output<LoDTensor>("Out");
      T* out_data = in_place ? out->data<T>() : out->mutable_data<T>(place);
      const int64_t numel = out->numel();
      if (!in_place) PADDLE_ENFORCE(cudaMemsetAsync(out_data, 0, sizeof(T) * numel, stream) == cudaSuccess);
      for (int i = in_place ? 1 : 0; i < N; ++i) {
        if (in_vars[i]->IsType<LoDTensor>()) {
          auto& in_t = in_vars[i]->Get<LoDTensor>();
          if (in_t.numel() == 0) continue;
          const int grid = static_cast<int>(std::min<int64_t>((numel + 255) / 256, 148 * 8));
          AddInto<T><<<grid, 256, 0, stream>>>(in_t.data<T>(), numel, out_data);
        } else if (in_vars[i]->IsType<SelectedRows>()) {
          if constexpr (std::is_floating_point<T>::value) {
            auto& in_t = in_vars[i]->Get<SelectedRows>();
            const b200::SparseGradView g = b200::ViewSparseGrad<T>(device, place, &in_t);
            if (g.n == 0) continue;
            PADDLE_ENFORCE_EQ(g.height, out->dims()[0]);
            b200::UpdateWorkspace w;
            b200::PrepareWorkspace(&w, g, sizeof(T), place, stream);
            B200EMB_CHECK(b200::Api<T>::add_to_tensor(g.rows, static_cast<const T*>(g.value), g.n, g.d, g.height,
                                                      out_data, w.ptr, w.bytes, w.flags, nullptr, stream));
            b200::SideTable::Instance().Release(device, g.plan);
          } else {
            PADDLE_THROW("sum: SelectedRows inputs need a float or double kernel");
          }
        } else {
          PADDLE_THROW("Variable type must be LoDTensor/SelectedRows.");
        }
      }
    } else if (out_var->IsType<SelectedRows>()) {
      // an in-place output would be overwritten while it is still an input: snapshot input 0 first
      std::unique_ptr<SelectedRows> in0;
      if (in_place) {
        auto& src = in_vars[0]->Get<SelectedRows>();
        const b200::SparseGradView g0 = b200::ViewSparseGrad<T>(device, place, &src);   // fused mode: device pointers
        const int64_t n0 = g0.n;
        in0.reset(new SelectedRows(std::vector<int64_t>(), src.height()));
        in0->mutable_rows()->resize(static_cast<size_t>(n0));
        in0->mutable_value()->Resize(src.value().dims());
        if (n0 > 0) {
          T* v = in0->mutable_value()->mutable_data<T>(place);
          int64_t* r = in0->mutable_rows()->CUDAMutableData(place);
          B200EMB_CHECK(b200emb_selected_rows_add_to(g0.rows, g0.value, n0, g0.d, sizeof(T), r, v, 0, stream));
        }
        b200::SideTable::Instance().Release(device, g0.plan);
      }
      auto get = [&](int i) -> const SelectedRows& {
        return (i == 0 && in0) ? *in0 : in_vars[i]->Get<SelectedRows>();
      };
      // fused mode: the inputs' rows / values are read from the device pointers the producers recorded
      std::vector<b200::SparseGradView> views(N);
      for (int i = 0; i < N; ++i) views[i] = b200::ViewSparseGrad<T>(device, place, &get(i));
      auto* out = context.Output<SelectedRows>("Out");
      int64_t total = 0;
      std::vector<int64_t> in_dim;
      for (int i = 0; i < N; ++i) {
        total += views[i].n;
        if (in_dim.empty() && views[i].n > 0) in_dim = framework::vectorize(get(i).value().dims());
      }
      if (in_dim.empty()) in_dim = framework::vectorize(get(N - 1).value().dims());   // all inputs empty
      else in_dim[0] = total;
      out->mutable_rows()->clear();
      out->mutable_rows()->resize(static_cast<size_t>(total));
      auto* out_value = out->mutable_value();
      out_value->Resize(framework::make_ddim(in_dim));
      T* out_data = out_value->mutable_data<T>(place);
      if (total == 0) return;
      const bool fused_out = b200::SideTable::Instance().fused();
      // fused mode keeps the concatenated rows in a device tensor of its own; rows() keeps only its size
      int64_t* out_rows = nullptr;
      std::shared_ptr<framework::Tensor> rows_dev;
      if (fused_out) {
        rows_dev = std::make_shared<framework::Tensor>();
        out_rows = rows_dev->mutable_data<int64_t>(framework::make_ddim({total}), place);
      } else {
        out_rows = out->mutable_rows()->CUDAMutableData(place);
      }
      const int64_t d = out_value->numel() / total;
      // Runtime InferShape (SumOp::InferShape, sum_op.cc:33-64, through RuntimeInferShapeContext::SetDim,
      // operator.cc:608-622) gives Out the complete dims of the inputs, i.e. their height, before the kernel
      // runs; the mini runner has no InferShape pass, so the same assignment happens here.
      for (int i = 0; i < N; ++i)
        if (views[i].n > 0) { out->set_height(views[i].height); break; }
      int64_t offset = 0;
      for (int i = 0; i < N; ++i) {
        const b200::SparseGradView& g = views[i];
        if (g.n == 0) continue;
        PADDLE_ENFORCE_EQ(out->height(), g.height);
        B200EMB_CHECK(b200emb_selected_rows_add_to(g.rows, g.value, g.n, d, sizeof(T), out_rows, out_data, offset,
                                                   stream));
        b200::SideTable::Instance().Release(device, g.plan);   // a plan of one input is of no use to the concatenation
        offset += g.n;
      }
      if (fused_out) {
        b200::SparseGradRecord rec;
        rec.ids = out_rows;
        rec.value = out_data;
        rec.n = total; rec.d = d; rec.height = out->height();
        rec.keep = rows_dev;
        b200::SideTable::Instance().Bind(device, out, rec);
      }
    } else {
      PADDLE_THROW("Unexpected branch, output variable type is %s", out_var->Type().name());
    }
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
// the reference registers four element types (operators/sum_op.cu:16-20)
REGISTER_OP_CUDA_KERNEL(sum, ops::SumB200Kernel<float>, ops::SumB200Kernel<double>, ops::SumB200Kernel<int>,
                        ops::SumB200Kernel<int64_t>);
