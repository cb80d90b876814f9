// segreduce_f64.cu — double instantiation of the fused duplicate-row reduction (segreduce.cuh).
#include "segreduce.cuh"

namespace b200emb {

template cudaError_t launch_update<double>(const SortedPlan&, const double*, int64_t, int64_t,
                                           int64_t, const UpdateArgs<double>&, bool, char*,
                                           const Workspace&, int32_t*, cudaStream_t);

}  // namespace b200emb
