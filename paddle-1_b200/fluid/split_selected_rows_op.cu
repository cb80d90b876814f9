// split_selected_rows_op.cu — replaces paddle/fluid/operators/split_selected_rows_op.{h,cu} of the reference
// (op `split_selected_rows`: X -> Out[]; attr height_sections; split_selected_rows_op.cc:20-88), whose GPU path
// copies the value rows with one cudaMemcpyAsync per row (split_selected_rows_op.h:83-99).  Here the section of
// every row, the stable partition, the rebased row ids and the value rows all come from one
// b200emb_split_selected_rows call; the op then slices the result into its outputs.
// Also `extract_rows` (extract_rows_op.cc:49-77): Out[n,1] = X.rows().
#include <vector>

#include "b200_op_common.h"

namespace paddle {
namespace operators {

namespace {
template <typename T> struct SplitApi;
template <> struct SplitApi<float> { static constexpr auto fn = b200emb_split_selected_rows_f32; };
template <> struct SplitApi<double> { static constexpr auto fn = b200emb_split_selected_rows_f64; };
}  // namespace

template <typename T>
class SplitSelectedRowsOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto* x = ctx.Input<framework::SelectedRows>("X");
    auto outs = ctx.MultiOutput<framework::SelectedRows>("Out");
    auto height_sections = ctx.Attr<std::vector<int>>("height_sections");
    PADDLE_ENFORCE_EQ(height_sections.size(), outs.size());
    auto place = ctx.GetPlace();
    auto stream = ctx.cuda_device_context().stream();
    const int64_t n = static_cast<int64_t>(x->rows().size());
    const int64_t ns = static_cast<int64_t>(outs.size());
    for (int64_t s = 0; s < ns; ++s) {
      outs[s]->set_height(height_sections[s]);              // split_selected_rows_op.h:74
      outs[s]->mutable_rows()->clear();
    }
    if (n == 0) return;
    const int64_t d = x->value().numel() / x->value().dims()[0];
    std::vector<int64_t> hs(height_sections.begin(), height_sections.end());
    framework::Tensor ws, rows_t, value_t, bounds_t;
    size_t ws_bytes = 0;
    void* wsp = b200::Workspace(&ws, n, d, sizeof(T), place, &ws_bytes);
    int64_t* out_rows = rows_t.mutable_data<int64_t>(framework::make_ddim({n}), place);
    T* out_value = value_t.mutable_data<T>(framework::make_ddim({n, d}), place);
    int64_t* bounds = bounds_t.mutable_data<int64_t>(framework::make_ddim({ns + 1}), place);
    B200EMB_CHECK(SplitApi<T>::fn(x->rows().CUDAData(place), x->value().data<T>(), n, d, hs.data(), ns, out_rows,
                                  out_value, bounds, wsp, ws_bytes, stream));
    std::vector<int64_t> hb(ns + 1), hrows(n);
    PADDLE_ENFORCE_EQ(cudaMemcpyAsync(hb.data(), bounds, sizeof(int64_t) * hb.size(), cudaMemcpyDeviceToHost, stream),
                      cudaSuccess);
    PADDLE_ENFORCE_EQ(cudaMemcpyAsync(hrows.data(), out_rows, sizeof(int64_t) * n, cudaMemcpyDeviceToHost, stream),
                      cudaSuccess);
    PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);
    for (int64_t s = 0; s < ns; ++s) {
      const int64_t cnt = hb[s + 1] - hb[s];
      if (cnt == 0) continue;                               // the reference leaves empty outputs unallocated (:75)
      auto dims = x->GetCompleteDims();
      dims[0] = cnt;
      T* dst = outs[s]->mutable_value()->mutable_data<T>(dims, place);
      outs[s]->mutable_rows()->assign(hrows.begin() + hb[s], hrows.begin() + hb[s + 1]);
      PADDLE_ENFORCE_EQ(cudaMemcpyAsync(dst, out_value + hb[s] * d, sizeof(T) * cnt * d, cudaMemcpyDeviceToDevice,
                                        stream), cudaSuccess);
    }
    PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);
  }
};

template <typename T>
class ExtractRowsOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto* x = ctx.Input<framework::SelectedRows>("X");
    auto* out = ctx.Output<framework::LoDTensor>("Out");
    const int64_t n = static_cast<int64_t>(x->rows().size());
    int64_t* dst = out->mutable_data<int64_t>(framework::make_ddim({n, 1}), ctx.GetPlace());   // extract_rows_op.cc:29-31
    B200EMB_CHECK(b200emb_extract_rows(x->rows().CUDAData(ctx.GetPlace()), n, dst, ctx.cuda_device_context().stream()));
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(split_selected_rows, ops::SplitSelectedRowsOpB200Kernel<float>,
                        ops::SplitSelectedRowsOpB200Kernel<double>);
REGISTER_OP_CUDA_KERNEL(extract_rows, ops::ExtractRowsOpB200Kernel<float>, ops::ExtractRowsOpB200Kernel<double>);
