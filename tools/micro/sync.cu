// Microbenchmark: __syncwarp with a runtime half-warp mask vs constant masks (sm_100a)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, int G) {
  __shared__ double sm[64];
  const int lane = threadIdx.x & 31, gw = lane / G;
  const unsigned gm = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (gw * G));
  sm[threadIdx.x & 63] = lane;
  double x = lane; long long t0, t1;
  __syncthreads();
  t0 = clock64();
  for (int i = 0; i < 256; i++) { sm[lane] = x; __syncwarp(gm); x = sm[lane ^ 1] + 1.0; __syncwarp(gm); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[0] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < 256; i++) { sm[lane] = x; __syncwarp(); x = sm[lane ^ 1] + 1.0; __syncwarp(); }
  t1 = clock64(); if (threadIdx.x == 0) cyc[1] = t1 - t0;
  t0 = clock64();
  if (gw == 0) { for (int i = 0; i < 256; i++) { sm[lane] = x; __syncwarp(0x0000ffffu); x = sm[lane ^ 1] + 1.0; __syncwarp(0x0000ffffu); } }
  else         { for (int i = 0; i < 256; i++) { sm[lane] = x; __syncwarp(0xffff0000u); x = sm[lane ^ 1] + 1.0; __syncwarp(0xffff0000u); } }
  t1 = clock64(); if (threadIdx.x == 0) cyc[2] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < 256; i++) { x = __shfl_xor_sync(gm, x, 1) + 1.0; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[3] = t1 - t0;
  t0 = clock64();
  for (int i = 0; i < 256; i++) { x = __shfl_xor_sync(0xffffffffu, x, 1) + 1.0; }
  t1 = clock64(); if (threadIdx.x == 0) cyc[4] = t1 - t0;
  out[threadIdx.x] = x;
}
int main() {
  double *o; long long *c, h[5];
  cudaMalloc(&o, 1024 * 8); cudaMalloc(&c, 5 * 8);
  k<<<1, 32>>>(o, c, 16); cudaMemcpy(h, c, 40, cudaMemcpyDeviceToHost);
  printf("per STS+sync+LDS+add+sync: runtime half mask %.1f | full mask %.1f | constant half masks (divergent halves) %.1f ; shfl+add runtime mask %.1f | full %.1f cycles\n",
         h[0] / 256.0, h[1] / 256.0, h[2] / 256.0, h[3] / 256.0, h[4] / 256.0);
  return 0;
}
