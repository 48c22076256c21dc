// Dependent-issue latency and issue rate of fp64 operations on one SM (B200: a dependent DFMA / DADD issues 8 cycles
// after its producer, independent ones every 2.2 cycles per scheduler) -- the figures profiles/r2_hw_tm.md reasons with.
// build + run: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o /tmp/fp64_latency scripts/fp64_latency.cu && /tmp/fp64_latency
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, double a, double b) {
  double x = out[threadIdx.x], y = x + 1.0, z = x + 2.0, w = x + 3.0;
  long long t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 4096; ++i) x = fma(x, a, b);
  long long t1 = clock64();
#pragma unroll 64
  for (int i = 0; i < 4096; ++i) x = x + b;
  long long t2 = clock64();
  // 2 independent chains
#pragma unroll 32
  for (int i = 0; i < 4096; ++i) { x = fma(x, a, b); y = fma(y, a, b); }
  long long t3 = clock64();
#pragma unroll 16
  for (int i = 0; i < 4096; ++i) { x = fma(x, a, b); y = fma(y, a, b); z = fma(z, a, b); w = fma(w, a, b); }
  long long t4 = clock64();
  float f = (float)x;
#pragma unroll 64
  for (int i = 0; i < 4096; ++i) { f = (float)((double)f + 1.0); }
  long long t5 = clock64();
  out[threadIdx.x] = x + y + z + w + f;
  if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; cyc[4] = t5 - t4; }
}
int main() {
  double *o; long long *c; cudaMalloc(&o, 1024 * 8); cudaMalloc(&c, 64); cudaMemset(o, 0, 1024 * 8);
  for (int warps : {1, 2, 4, 8}) {
    k<<<1, 32 * warps>>>(o, c, 0.999, 1e-3); cudaDeviceSynchronize();
    long long h[5]; cudaMemcpy(h, c, 40, cudaMemcpyDeviceToHost);
    printf("warps/SM %d (per SMSP %g): dep DFMA %.1f  dep DADD %.1f  2 chains %.1f per pair  4 chains %.1f per quad  F2F+DADD+F2F %.1f cycles\n", warps, warps / 4.0, h[0] / 4096.0, h[1] / 4096.0, h[2] / 4096.0, h[3] / 4096.0, h[4] / 4096.0);
  }
  return 0;
}
