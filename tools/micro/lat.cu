// Microbenchmark: dependent-issue latencies on sm_100a (DFMA, DADD, LDS.64, LDS.128 broadcast, SHFL, STS->sync->LDS)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, double a, double b) {
  __shared__ double sm[256];
  sm[threadIdx.x] = threadIdx.x * 0.5;
  __syncthreads();
  double x = a; long long t0, t1;
  t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 1024; i++) x = fma(x, b, a);
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
  t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 1024; i++) x = x + b;
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
  int idx = (int)x & 0; 
  t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 1024; i++) idx = (int)sm[idx + (threadIdx.x & 1)] & 0 ? 1 : (int)sm[idx] & 1 ? 0 : idx;  // dependent LDS + cvt
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
  t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 1024; i++) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31);
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < 256; i++) { sm[threadIdx.x] = x; __syncwarp(); x = sm[(threadIdx.x + 1) & 31] + 1.0; __syncwarp(); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
  // 4 independent DFMA chains (ILP)
  double y0 = a, y1 = b, y2 = a + 1, y3 = b + 1;
  t0 = clock64();
#pragma unroll 64
  for (int i = 0; i < 1024; i++) { y0 = fma(y0, b, a); y1 = fma(y1, b, a); y2 = fma(y2, b, a); y3 = fma(y3, b, a); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[5] = t1 - t0;
  out[threadIdx.x] = x + idx + y0 + y1 + y2 + y3;
}
int main() {
  double *o; long long *c, h[6];
  cudaMalloc(&o, 256 * 8); cudaMalloc(&c, 6 * 8);
  for (int th : {32, 16 * 32}) {
    k<<<1, th>>>(o, c, 1.0000001, 0.9999999); cudaMemcpy(h, c, 48, cudaMemcpyDeviceToHost);
    printf("threads %d: DFMA dep %.1f  DADD dep %.1f  LDS-chain %.1f  SHFL.64 dep %.1f  STS-sync-LDS-add-sync %.1f  4xDFMA ILP per-4 %.1f cycles\n", th,
           h[0] / 1024.0, h[1] / 1024.0, h[2] / 1024.0, h[3] / 1024.0, h[4] / 256.0, h[5] / 1024.0);
  }
  return 0;
}
