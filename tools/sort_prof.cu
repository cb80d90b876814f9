// Developer tool: times the sort plan and the head scan alone and checks them against a host stable sort.
//   nvcc -O3 -std=c++20 -gencode arch=compute_100a,code=sm_100a -DB200EMB_PROFILE -o tools/sort_prof tools/sort_prof.cu
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <numeric>
#include <vector>
#include "../paddle-1_b200/csrc/sort.cu"
namespace b200emb {
bool cluster_sort_fits(int64_t) { return false; }
cudaError_t launch_cluster_sort(const int64_t*, int64_t, int64_t, int, uint32_t*, uint32_t*, int32_t*, cudaStream_t) { return cudaErrorNotSupported; }
}
using namespace b200emb;
int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 106496;
  const int64_t height = argc > 2 ? atoll(argv[2]) : 1000000;
  std::vector<int64_t> h(n);
  srand(1);
  for (int i = 0; i < n; ++i) h[i] = (rand() % 8 == 0) ? 12345 : (static_cast<int64_t>(rand()) * 7919) % height;
  Workspace L = make_workspace(n, 64, 4);
  int64_t* d_rows; char* ws;
  cudaMalloc(&d_rows, n * 8); cudaMalloc(&ws, L.total);
  cudaMemcpy(d_rows, h.data(), n * 8, cudaMemcpyHostToDevice);
  SortedPlan plan;
  for (int it = 0; it < 3; ++it) launch_plan(d_rows, n, height, ws, L, nullptr, 0, &plan);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int it = 0; it < 20; ++it) launch_plan(d_rows, n, height, ws, L, nullptr, 0, &plan);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("n=%d passes=%d  %.2f us per plan (%s)\n", n, plan.passes, ms / 20 * 1e3, cudaGetErrorString(cudaGetLastError()));
  cudaEventRecord(e0);
  for (int it = 0; it < 20; ++it)
    launch_head_scan(plan, n, static_cast<uint32_t>(height), ws, L, nullptr, nullptr, 0);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  cudaEventElapsedTime(&ms, e0, e1);
  printf("head scan %.2f us (%s)\n", ms / 20 * 1e3, cudaGetErrorString(cudaGetLastError()));
  {  // check against a host stable sort
    std::vector<uint32_t> k(n), pm(n), ui(n), want(n);
    cudaMemcpy(k.data(), plan.keys, n * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(pm.data(), plan.perm, n * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(ui.data(), ws + L.off_uidx, n * 4, cudaMemcpyDeviceToHost);
    std::iota(want.begin(), want.end(), 0u);
    std::stable_sort(want.begin(), want.end(), [&](uint32_t a, uint32_t b) { return h[a] < h[b]; });
    long bad = 0; uint32_t u = 0;
    for (int i = 0; i < n; ++i) {
      if (i > 0 && h[want[i]] != h[want[i - 1]]) ++u;
      bad += pm[i] != want[i] || k[i] != static_cast<uint32_t>(h[want[i]]) || ui[i] != u;
    }
    PlanHeader hd; cudaMemcpy(&hd, ws + L.off_plan, sizeof(hd), cudaMemcpyDeviceToHost);
    printf("check: %ld mismatches, U=%u (want %u)\n", bad, (unsigned)hd.n_unique, u + 1);
  }
  return 0;
}
