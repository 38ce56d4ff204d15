// Microbenchmark: do LDS.64 and SHFL share one issue/data path on sm_100a?  Throughput (not latency) per SM of
//   A  8 lane-distinct LDS.64 per iteration and warp
//   B  8 64-bit shuffles (16 SHFL.32) per iteration and warp
//   C  A and B interleaved
//   D  A plus 16 independent DFMA
// with 16 warps per SM so that latencies are hidden.  If C costs max(A, B) the two run on separate pipes and the
// vector operand of the small mat-vecs could move from shared memory to shuffles; if C = A + B it cannot.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double lds64(unsigned a) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ double shfl64(double v, int src) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  asm volatile("shfl.sync.idx.b32 %0, %0, %1, 0x1f, 0xffffffff;" : "+r"(lo) : "r"(src));
  asm volatile("shfl.sync.idx.b32 %0, %0, %1, 0x1f, 0xffffffff;" : "+r"(hi) : "r"(src));
  return __hiloint2double(hi, lo);
}

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(double *out, long long *cyc, int iters, double a, double b) {
  __shared__ double sm[512 + 64];
  sm[threadIdx.x] = threadIdx.x * 0.5;
  if (threadIdx.x < 64) sm[512 + threadIdx.x] = 1.0;
  __syncthreads();
  const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + 8u * threadIdx.x;
  const int src = (threadIdx.x + 1) & 31;
  double acc[8] = {a, b, a, b, a, b, a, b};
  double f[16];
#pragma unroll
  for (int j = 0; j < 16; j++) f[j] = a + j;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
      if (MODE == 0 || MODE == 2 || MODE == 3) acc[j] += lds64(base + 8u * j + 8u * (i & 7));
      if (MODE == 1 || MODE == 2) acc[j] = shfl64(acc[j], src);
      if (MODE == 3) { f[2 * j] = fma(f[2 * j], b, a); f[2 * j + 1] = fma(f[2 * j + 1], b, a); }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[MODE] = t1 - t0;
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) s += acc[j];
#pragma unroll
  for (int j = 0; j < 16; j++) s += f[j];
  out[blockIdx.x * 512 + threadIdx.x] = s;
}

int main() {
  double *o;
  long long *c, h[4];
  cudaMalloc(&o, 148 * 512 * 8);
  cudaMalloc(&c, 4 * 8);
  const int iters = 4096;
  for (int rep = 0; rep < 2; rep++) {
    k<0><<<148, 512>>>(o, c, iters, 1.0000001, 0.9999999);
    k<1><<<148, 512>>>(o, c, iters, 1.0000001, 0.9999999);
    k<2><<<148, 512>>>(o, c, iters, 1.0000001, 0.9999999);
    k<3><<<148, 512>>>(o, c, iters, 1.0000001, 0.9999999);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(h, c, 32, cudaMemcpyDeviceToHost);
  // per SM and iteration: 16 warps x 8 ops
  printf("cycles per warp-level op group (16 warps/SM): 8xLDS.64 %.2f | 8xSHFL.64 %.2f | both %.2f | 8xLDS.64+16xDFMA %.2f  (per warp-iteration, SM-wide)\n",
         h[0] / (double)iters / 16.0, h[1] / (double)iters / 16.0, h[2] / (double)iters / 16.0, h[3] / (double)iters / 16.0);
  printf("=> LDS.64 %.2f cycles/warp-instruction, SHFL.64 %.2f, LDS.64+SHFL.64 %.2f\n", h[0] / (double)iters / 128.0,
         h[1] / (double)iters / 128.0, h[2] / (double)iters / 128.0);
  return 0;
}
