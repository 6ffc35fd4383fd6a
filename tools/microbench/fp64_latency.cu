// Dependent-issue latency of the FP64 pipe, 64-bit shuffles and shared-memory loads on this GPU (one warp, clock64).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, double a, double b) {
  __shared__ double sm[64];
  sm[threadIdx.x] = a; sm[threadIdx.x + 32] = b;
  __syncthreads();
  double x = a;
  long long t0, t1;
  const int N = 4096;
  t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = __dmul_rn(x, b);
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = (t1 - t0); out[0] = x;
  x = a; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = __dadd_rn(x, b);
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = (t1 - t0); out[1] = x;
  x = a; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = fma(x, b, a);
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = (t1 - t0); out[2] = x;
  x = a; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) x = __shfl_up_sync(0xffffffffu, x, 1);
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = (t1 - t0); out[3] = x;
  float xf = (float)a, bf = (float)b; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) xf = fmaf(xf, bf, bf);
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = (t1 - t0); out[4] = xf;
  int idx = threadIdx.x; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) idx = (int)sm[idx & 63] & 31;
  t1 = clock64(); if (threadIdx.x == 0) cyc[5] = (t1 - t0); out[5] = idx;
  // two independent FP64 chains: does the pipe overlap them?
  double y = b; x = a; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) { x = fma(x, b, a); y = fma(y, a, b); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[6] = (t1 - t0); out[6] = x + y;
  double z = a + b, u = a - b; x = a; y = b; t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) { x = fma(x, b, a); y = fma(y, a, b); z = fma(z, a, b); u = fma(u, b, a); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[7] = (t1 - t0); out[7] = x + y + z + u;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 64); cudaMalloc(&cyc, 64);
  for (int nw = 1; nw <= 4; nw *= 2) {
    k<<<1, 32 * nw>>>(out, cyc, 1.0000001, 0.9999999);
    cudaDeviceSynchronize();
    long long h[8]; cudaMemcpy(h, cyc, 64, cudaMemcpyDeviceToHost);
    printf("warps/block %d: cycles per dependent op: DMUL %.1f DADD %.1f DFMA %.1f SHFL64 %.1f FFMA %.1f LDS.64+cvt %.1f | 2 chains DFMA %.1f 4 chains %.1f (per loop trip)\n",
           nw, h[0] / 4096.0, h[1] / 4096.0, h[2] / 4096.0, h[3] / 4096.0, h[4] / 4096.0, h[5] / 4096.0, h[6] / 4096.0, h[7] / 4096.0);
  }
  return 0;
}
