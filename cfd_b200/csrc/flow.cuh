// flow.cuh -- device helpers shared by the flag-ordered sweep kernels (tdma_flow_kernels.cu, tdma_cluster_kernels.cu).
#pragma once
#include "common.cuh"
#include "tma.cuh"

namespace cfd {

__device__ __forceinline__ double rcp_nr(double x) {   // == rcp() of tdma_kernels.cu
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(fma(-x, r, 1.0), r, r);
  r = fma(fma(-x, r, 1.0), r, r);
  return r;
}

__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Poll until *p >= want (wrap-safe) or about two seconds have passed; false = timed out (abort flag raised).
// SYS: the flag may be written by another GPU.
template <bool SYS = true>
__device__ __forceinline__ bool wait_flag(const int* p, int want, int* abort_flag) {
  if ((SYS ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - want >= 0) return true;
  const long long t0 = clock64();
  for (;;) {
    __nanosleep(40);
    if ((SYS ? ld_acquire_sys(p) : ld_acquire_gpu(p)) - want >= 0) return true;
    if (clock64() - t0 > 4000000000LL || *(volatile int*)abort_flag) {   // ~2 s at 2 GHz
      atomicExch(abort_flag, 1);
      return false;
    }
  }
}

// position in the per-block work order -> line of the colour: slab-edge lines (first, last) come first
__device__ __forceinline__ int flow_line(int idx, int n) { return idx == 0 ? 0 : (idx == 1 ? n - 1 : idx - 1); }

// A block's work items in order: (pass, position in the pass).  Every role of the block walks the same sequence.
struct FlowIter {
  int p, idx, first, step;   // first / step: this block's (or cluster's) number and the number of blocks (clusters)
  __device__ __forceinline__ void skip(const FlowParams& P) {
    while (p < P.npass && idx >= P.nlines[(P.colour0 + p) & 1]) ++p;
  }
  __device__ __forceinline__ void init(const FlowParams& P, int first_ = (int)blockIdx.x, int step_ = (int)gridDim.x) {
    p = 0;
    first = first_;
    step = step_;
    idx = first;
    skip(P);
  }
  __device__ __forceinline__ bool done(const FlowParams& P) const { return p >= P.npass; }
  __device__ __forceinline__ int colour(const FlowParams& P) const { return (P.colour0 + p) & 1; }
  __device__ __forceinline__ int line(const FlowParams& P) const {
    const int c = (P.colour0 + p) & 1;
    return P.i_first[c] + 2 * flow_line(idx, P.nlines[c]);
  }
  __device__ __forceinline__ void next(const FlowParams& P) {
    idx += step;
    if (idx >= P.nlines[(P.colour0 + p) & 1]) {
      idx = first;
      ++p;
      skip(P);
    }
  }
};

__device__ __forceinline__ int lds_acquire(unsigned addr) {
  int v;
  asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_release(unsigned addr, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

}  // namespace cfd
