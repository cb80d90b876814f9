// split_ids_op.cu / merge_ids — CUDA kernels for two ops that are CPU-only in the reference
// (`PADDLE_THROW("SplitIds do not support GPU kernel")`, split_ids_op.h:30-32; merge_ids_op.h:30-32).
//   split_ids  {Ids -> Out[]}      LoDTensor Ids: shard g gets the ids with id % shard_num == g, input order and
//                                  duplicates kept (split_ids_op.h:36-58) — b200emb_split_ids on the device.
//                                  SelectedRows Ids (:59-92): the row ids live in host memory (Vector<int64_t>), so
//                                  they are partitioned on the host exactly like the reference does, INCLUDING its
//                                  id_to_index map (a duplicated row id takes the value of its LAST occurrence);
//                                  the value rows move with one device gather.
//   merge_ids  {Ids, X[] -> Out}   merge_ids_op.h:63-87 — b200emb_merge_ids_* after a split of the same ids.
#include <unordered_map>
#include <vector>

#include "b200_op_common.h"

namespace paddle {
namespace operators {

namespace {
constexpr int64_t kIdSpace = 0xFFFFFFFELL;   // ids are sorted as 32-bit keys inside the library

template <typename T> struct GatherApi;      // "lookup_table" over raw 4- and 8-byte elements
template <> struct GatherApi<float> { static constexpr auto fn = b200emb_lookup_table_f32; using U = float; };
template <> struct GatherApi<int64_t> { static constexpr auto fn = b200emb_lookup_table_f64; using U = double; };
template <typename T> struct MergeApi;
template <> struct MergeApi<float> { static constexpr auto fn = b200emb_merge_ids_f32; };
template <> struct MergeApi<double> { static constexpr auto fn = b200emb_merge_ids_f64; };

void CheckStatus(const framework::Tensor& status, cudaStream_t stream, const char* what) {
  int32_t h = 0;
  PADDLE_ENFORCE_EQ(cudaMemcpyAsync(&h, status.data<int32_t>(), sizeof(h), cudaMemcpyDeviceToHost, stream),
                    cudaSuccess);
  PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);
  PADDLE_ENFORCE_EQ(h, 0, "%s: an id is negative or does not fit 32 bits", what);
}
}  // namespace

template <typename T>
class SplitIdsOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto place = ctx.GetPlace();
    auto stream = ctx.cuda_device_context().stream();
    const auto* ids_var = ctx.InputVar("Ids");
    if (ids_var->IsType<framework::LoDTensor>()) {
      const auto* ids_t = ctx.Input<framework::LoDTensor>("Ids");
      auto outs = ctx.MultiOutput<framework::LoDTensor>("Out");
      const int64_t shard_num = static_cast<int64_t>(outs.size());
      const int64_t n = ids_t->dims()[0];
      PADDLE_ENFORCE((std::is_same<T, int64_t>::value), "ids are int64 (split_ids_op.cc:75-77)");
      const int64_t* ids = reinterpret_cast<const int64_t*>(ids_t->data<T>());
      framework::Tensor ws, grouped, bounds, status;
      size_t ws_bytes = 0;
      void* wsp = b200::Workspace(&ws, std::max<int64_t>(n, 1), 1, 4, place, &ws_bytes);
      int64_t* g = grouped.mutable_data<int64_t>(framework::make_ddim({std::max<int64_t>(n, 1)}), place);
      int64_t* b = bounds.mutable_data<int64_t>(framework::make_ddim({shard_num + 1}), place);
      int32_t* st = status.mutable_data<int32_t>(framework::make_ddim({1}), place);
      PADDLE_ENFORCE_EQ(cudaMemsetAsync(st, 0, sizeof(int32_t), stream), cudaSuccess);
      B200EMB_CHECK(b200emb_split_ids(ids, n, kIdSpace, shard_num, g, b, wsp, ws_bytes, st, stream));
      std::vector<int64_t> hb(shard_num + 1);
      PADDLE_ENFORCE_EQ(cudaMemcpyAsync(hb.data(), b, sizeof(int64_t) * hb.size(), cudaMemcpyDeviceToHost, stream),
                        cudaSuccess);
      CheckStatus(status, stream, "split_ids");
      for (int64_t s = 0; s < shard_num; ++s) {
        const int64_t cnt = hb[s + 1] - hb[s];
        auto* dst = outs[s]->mutable_data<T>(framework::make_ddim({cnt, 1}), place);
        if (cnt > 0)
          PADDLE_ENFORCE_EQ(cudaMemcpyAsync(dst, g + hb[s], sizeof(int64_t) * cnt, cudaMemcpyDeviceToDevice, stream),
                            cudaSuccess);
      }
      PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);   // `grouped` dies with this scope
    } else if (ids_var->IsType<framework::SelectedRows>()) {
      const auto* in = ctx.Input<framework::SelectedRows>("Ids");
      const auto& dims = in->value().dims();
      const auto& rows = in->rows();
      PADDLE_ENFORCE_EQ(dims[0], static_cast<int64_t>(rows.size()), "");
      auto outs = ctx.MultiOutput<framework::SelectedRows>("Out");
      const size_t shard_num = outs.size();
      for (auto* out : outs) out->mutable_rows()->clear();
      std::unordered_map<int64_t, size_t> id_to_index;                 // split_ids_op.h:71-76
      for (size_t i = 0; i < rows.size(); ++i) {
        id_to_index[rows[i]] = i;
        outs[static_cast<size_t>(rows[i]) % shard_num]->mutable_rows()->push_back(rows[i]);
      }
      const int64_t width = dims[1], n = static_cast<int64_t>(rows.size());
      std::vector<int64_t> src;                                        // source row of every output row
      src.reserve(rows.size());
      for (auto* out : outs)
        for (int64_t r : out->rows()) src.push_back(static_cast<int64_t>(id_to_index[r]));
      framework::Tensor src_d, all;
      using U = typename GatherApi<T>::U;
      U* gathered = nullptr;
      if (n > 0) {
        int64_t* sd = src_d.mutable_data<int64_t>(framework::make_ddim({n}), place);
        PADDLE_ENFORCE_EQ(cudaMemcpyAsync(sd, src.data(), sizeof(int64_t) * n, cudaMemcpyHostToDevice, stream),
                          cudaSuccess);
        gathered = reinterpret_cast<U*>(all.mutable_data<T>(framework::make_ddim({n, width}), place));
        B200EMB_CHECK(GatherApi<T>::fn(reinterpret_cast<const U*>(in->value().data<T>()), n, width, sd, n, -1,
                                       gathered, nullptr, stream));
      }
      int64_t off = 0;
      for (auto* out : outs) {
        out->set_height(in->height());
        const int64_t cnt = static_cast<int64_t>(out->rows().size());
        T* dst = out->mutable_value()->mutable_data<T>(framework::make_ddim({cnt, width}), place);
        if (cnt > 0)
          PADDLE_ENFORCE_EQ(cudaMemcpyAsync(dst, gathered + off * width, sizeof(T) * cnt * width,
                                            cudaMemcpyDeviceToDevice, stream), cudaSuccess);
        off += cnt;
      }
      PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);
    } else {
      PADDLE_THROW("Ids should be LoDTensor or SelectedRows");
    }
  }
};

template <typename T>
class MergeIdsOpB200Kernel : public framework::OpKernel<T> {
 public:
  void Compute(const framework::ExecutionContext& ctx) const override {
    auto place = ctx.GetPlace();
    auto stream = ctx.cuda_device_context().stream();
    const auto* ids_var = ctx.InputVar("Ids");
    PADDLE_ENFORCE(ids_var->IsType<framework::LoDTensor>(), "only support to merge Ids of LoDTensor");
    const auto& ids_t = ids_var->Get<framework::LoDTensor>();
    const int64_t n = ids_t.dims()[0];
    auto xs = ctx.MultiInput<framework::LoDTensor>("X");
    auto* out = ctx.Output<framework::LoDTensor>("Out");
    int64_t batch = 0, emb = 0;
    for (auto* x : xs)
      if (framework::product(x->dims()) != 0) {
        if (emb == 0) emb = x->dims()[1];
        PADDLE_ENFORCE_EQ(emb, x->dims()[1], "embedding size of all input should be the same");
        batch += x->dims()[0];
      }
    PADDLE_ENFORCE_EQ(batch, n, "the batch size of ids and merged embedding value should be the same");
    const int64_t shard_num = static_cast<int64_t>(xs.size());
    T* dst = out->mutable_data<T>(framework::make_ddim({batch, emb}), place);
    if (n == 0) return;
    // the permutation of a split of the same ids, then the shards' blocks laid out shard after shard
    framework::Tensor ws, grouped, bounds, status, cat;
    size_t ws_bytes = 0;
    void* wsp = b200::Workspace(&ws, n, emb, sizeof(T), place, &ws_bytes);
    int64_t* g = grouped.mutable_data<int64_t>(framework::make_ddim({n}), place);
    int64_t* b = bounds.mutable_data<int64_t>(framework::make_ddim({shard_num + 1}), place);
    int32_t* st = status.mutable_data<int32_t>(framework::make_ddim({1}), place);
    PADDLE_ENFORCE_EQ(cudaMemsetAsync(st, 0, sizeof(int32_t), stream), cudaSuccess);
    B200EMB_CHECK(b200emb_split_ids(ids_t.data<int64_t>(), n, kIdSpace, shard_num, g, b, wsp, ws_bytes, st, stream));
    std::vector<int64_t> hb(shard_num + 1);
    PADDLE_ENFORCE_EQ(cudaMemcpyAsync(hb.data(), b, sizeof(int64_t) * hb.size(), cudaMemcpyDeviceToHost, stream),
                      cudaSuccess);
    CheckStatus(status, stream, "merge_ids");
    T* c = cat.mutable_data<T>(framework::make_ddim({n, emb}), place);
    for (int64_t s = 0; s < shard_num; ++s) {
      const int64_t rows = framework::product(xs[s]->dims()) != 0 ? xs[s]->dims()[0] : 0;
      PADDLE_ENFORCE_EQ(hb[s + 1] - hb[s], rows, "after merge, all data in x_tensor should be used");
      if (rows > 0)
        PADDLE_ENFORCE_EQ(cudaMemcpyAsync(c + hb[s] * emb, xs[s]->data<T>(), sizeof(T) * rows * emb,
                                          cudaMemcpyDeviceToDevice, stream), cudaSuccess);
    }
    B200EMB_CHECK(MergeApi<T>::fn(c, n, emb, shard_num, wsp, ws_bytes, dst, stream));
    PADDLE_ENFORCE_EQ(cudaStreamSynchronize(stream), cudaSuccess);
  }
};

}  // namespace operators
}  // namespace paddle

namespace ops = paddle::operators;
REGISTER_OP_CUDA_KERNEL(split_ids, ops::SplitIdsOpB200Kernel<int64_t>, ops::SplitIdsOpB200Kernel<float>);
REGISTER_OP_CUDA_KERNEL(merge_ids, ops::MergeIdsOpB200Kernel<float>, ops::MergeIdsOpB200Kernel<double>);
