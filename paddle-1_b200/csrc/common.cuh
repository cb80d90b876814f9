// common.cuh — shared device helpers for the sm_100a sparse-embedding kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>

namespace b200emb {

constexpr uint32_t kNoKey = 0xFFFFFFFFu;  // never a valid sort key (height < 2^32-1)
constexpr int kBlock = 256;               // threads per CTA for the row kernels

// ---------------------------------------------------------------------------
// IEEE round-to-nearest arithmetic that the compiler may not contract into FMAs: the
// reference CPU build has no FMA (cmake/simd.cmake:11), so parity needs separate roundings.
template <typename T> struct Arith;
template <> struct Arith<float> {
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
};
template <> struct Arith<double> {
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
  static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
};

// ---------------------------------------------------------------------------
// A row of D elements is spread over a "group" of LPR lanes; each lane owns NV vectors of
// EPV elements (16-byte vectors when VEC, single elements otherwise):
//   column(l, k) = col0 + (l + k*LPR) * EPV,   k < NV.
// A pass covers LPR*NV*EPV columns; wider rows loop over col0.
template <typename T, bool VEC> struct Epv { static constexpr int value = VEC ? int(16 / sizeof(T)) : 1; };

template <typename T, int EPV> struct Pack { T v[EPV]; };

template <typename T, bool VEC>
__device__ __forceinline__ void load_vec(const T* __restrict__ p, T (&dst)[Epv<T, VEC>::value]) {
  if constexpr (VEC) {
    const int4 raw = *reinterpret_cast<const int4*>(p);
    *reinterpret_cast<int4*>(dst) = raw;
  } else {
    dst[0] = *p;
  }
}
// streaming read of data touched once (gradient rows): bypass L1 allocation
template <typename T, bool VEC>
__device__ __forceinline__ void load_vec_stream(const T* __restrict__ p, T (&dst)[Epv<T, VEC>::value]) {
  if constexpr (VEC) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    *reinterpret_cast<int4*>(dst) = r;
  } else {
    dst[0] = __ldg(p);
  }
}
// read of data another CTA of the same grid has just published (after a fence): L2 only
template <typename T, bool VEC>
__device__ __forceinline__ void load_vec_cg(const T* p, T (&dst)[Epv<T, VEC>::value]) {
  if constexpr (VEC) {
    const int4 raw = __ldcg(reinterpret_cast<const int4*>(p));
    *reinterpret_cast<int4*>(dst) = raw;
  } else {
    dst[0] = __ldcg(p);
  }
}
template <typename T, bool VEC>
__device__ __forceinline__ void store_vec(T* p, const T (&src)[Epv<T, VEC>::value]) {
  if constexpr (VEC) {
    *reinterpret_cast<int4*>(p) = *reinterpret_cast<const int4*>(src);
  } else {
    *p = src[0];
  }
}
template <typename T, bool VEC>
__device__ __forceinline__ void store_vec_stream(T* p, const T (&src)[Epv<T, VEC>::value]) {
  if constexpr (VEC) {
    const int4 r = *reinterpret_cast<const int4*>(src);
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
  } else {
    *p = src[0];
  }
}

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Lanes of the warp that hold the same BITS-bit digit (valid lanes only).  Built from BITS warp
// ballots: the hardware MATCH.ANY instruction measured ~1 us per warp-instruction on sm_100a
// (profiles/r1_notes.md), the ballot form is ~20 issue slots.
template <int BITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t digit, bool valid) {
  uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    const bool bit = (digit >> b) & 1u;
    const uint32_t m = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// ---------------------------------------------------------------------------
// Consumer side of the sharded path's arrival flags (WaitSpec, internal.h); call with the whole CTA.
__device__ __forceinline__ void wait_for_flags(const unsigned long long* flags, const unsigned long long* epoch,
                                               int world, int32_t* status) {
  if (flags == nullptr) return;
  if (static_cast<int>(threadIdx.x) < world) {
    const unsigned long long want = *reinterpret_cast<const volatile unsigned long long*>(epoch);
    unsigned long long t0, t, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + threadIdx.x) : "memory");
      if (v >= want) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t - t0 > 20000000000ull) {   // 20 s: a rank is gone; do not hang the device
        if (status) atomicOr(status, 4);
        break;
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
}

// Producer side (SignalSpec, internal.h); call with the whole CTA as the last thing a kernel does.
__device__ __forceinline__ void signal_peers(unsigned long long* const* flag_dst, int64_t* const* payload_dst,
                                             const int64_t* starts, const unsigned long long* epoch,
                                             unsigned int* done_ctr, int world) {
  if (flag_dst == nullptr) return;
  __syncthreads();                   // every thread's stores to peer memory precede the barrier ...
  __shared__ bool s_last_cta;
  if (threadIdx.x == 0) {
    __threadfence();                 // ... one cumulative gpu-scope fence per CTA orders them before its arrival
    s_last_cta = atomicAdd(done_ctr, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last_cta) return;
  // The last CTA has observed every arrival; its release at system scope below is cumulative over everything that
  // happened before those arrivals (PTX memory model: causality order is transitive across scopes), so the one
  // system-scope fence of the whole kernel is the one inside st.release.sys — a system fence costs microseconds
  // here, and it used to be paid three times in a row.
  if (static_cast<int>(threadIdx.x) < world) {
    const int r = threadIdx.x;
    if (payload_dst) {
      payload_dst[r][0] = starts[r];
      payload_dst[r][1] = starts[r + 1] - starts[r];
    }
    const unsigned long long e = *reinterpret_cast<const volatile unsigned long long*>(epoch);
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag_dst[r]), "l"(e) : "memory");
  }
  if (threadIdx.x == 0) *done_ctr = 0;
}

// ---------------------------------------------------------------------------
// Row-shape dispatch: picks (LPR, NV, VEC) from D and pointer alignment and calls
// f.template operator()<LPR, NV, VEC>().
template <typename T, class F>
inline void dispatch_row_shape(int64_t d, bool aligned16, F&& f) {
  constexpr int epv = int(16 / sizeof(T));
  if (aligned16 && d % epv == 0) {
    const int64_t vecs = d / epv;
    if (vecs <= 8) f.template operator()<8, 1, true>();
    else if (vecs <= 16) f.template operator()<16, 1, true>();
    else if (vecs <= 32) f.template operator()<32, 1, true>();
    else if (vecs <= 64) f.template operator()<32, 2, true>();
    else f.template operator()<32, 4, true>();
  } else {
    if (d <= 8) f.template operator()<8, 1, false>();
    else if (d <= 16) f.template operator()<16, 1, false>();
    else f.template operator()<32, 1, false>();
  }
}

inline bool is_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------
// Programmatic dependent launch (opt-in, B200EMB_PDL=1): a row kernel launched right behind another kernel of
// the same stream may be set up (launch processing, CTA placement as SMs drain) while its predecessor is
// still running; it blocks in pdl_wait() — the first thing it executes, before any global memory access —
// until the predecessor has completed and its writes are visible.  Without the launch attribute both
// instructions are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
  static const bool on = [] {
    const char* v = std::getenv("B200EMB_PDL");
    return v != nullptr && v[0] != '0';
  }();
  return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_row_kernel(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t s,
                                     Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(static_cast<unsigned>(block));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  if (pdl_enabled()) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace b200emb
