// segreduce_f32.cu — float instantiation of the fused duplicate-row reduction (segreduce.cuh).
#include "segreduce.cuh"

namespace b200emb {

int choose_chunk(int64_t n, int groups_per_cta, int ctas_per_sm) {
  // one full wave of lane groups across the chip's 148 SMs when that fits a chunk of [4, 64]; batches that need
  // more than 8 waves at 64 take up to 128 while a CTA stays within 1024 elements (fewer, longer CTAs: C5's update
  // 1372 -> 1296 us; C3, 3 waves, is faster at 64).  A function of (N, D) alone, like everything about the geometry.
  const int64_t target_groups = 148LL * groups_per_cta * ctas_per_sm;
  int64_t c = ceil_div(n, target_groups);
  const int64_t cap = c > 8 * 64 ? std::max<int64_t>(kMaxChunk / 2, std::min<int64_t>(kMaxChunk, 1024 / groups_per_cta))
                                 : kMaxChunk / 2;
  c = std::max<int64_t>(kMinChunk, std::min<int64_t>(cap, c));
  return static_cast<int>(c);
}

template cudaError_t launch_update<float>(const SortedPlan&, const float*, int64_t, int64_t, int64_t,
                                          const UpdateArgs<float>&, bool, char*, const Workspace&,
                                          int32_t*, cudaStream_t);

}  // namespace b200emb

#ifdef B200EMB_TRACE
// developer builds only: copy the per-CTA phase stamps of the last float k_chunk_reduce launch to the host
extern "C" int b200emb_debug_trace(unsigned long long* host_out, size_t bytes) {
  return static_cast<int>(cudaMemcpyFromSymbol(host_out, b200emb::g_trace, bytes));
}
extern "C" int b200emb_debug_trace_clear() {
  void* p = nullptr;
  if (cudaGetSymbolAddress(&p, b200emb::g_trace) != cudaSuccess) return -1;
  return static_cast<int>(cudaMemset(p, 0, sizeof(b200emb::g_trace)));
}
#endif
