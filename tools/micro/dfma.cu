// dfma.cu — FP64 FMA throughput of the box (the denominator of every "fraction of FP64 peak" in this repo).
// Each thread runs 8 independent DFMA chains (registers only); the grid fills every SM with 2048 threads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dfma dfma.cu && ./dfma
// Prints GFLOP/s (2 flop per DFMA) for a few occupancies, best of 5, CUDA events.
#include <cstdio>
#include <cuda_runtime.h>

template <int CH>
__global__ void k_dfma(double *out, int iters, double a, double b) {
  double x[CH];
#pragma unroll
  for (int c = 0; c < CH; c++) x[c] = threadIdx.x * 1e-3 + c;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
#pragma unroll
      for (int c = 0; c < CH; c++) x[c] = fma(x[c], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int c = 0; c < CH; c++) s += x[c];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  double *out;
  cudaMalloc(&out, sizeof(double) * sms * 2048);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 4096;
  constexpr int CH = 8;
  double best_all = 0;
  for (int tpb : {128, 256, 512, 1024}) {
    for (int bps : {1, 2, 4, 8}) {
      if (tpb * bps > 2048) continue;
      double best = 0;
      for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k_dfma<CH><<<sms * bps, tpb>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fl = 2.0 * CH * 8 * (double)iters * tpb * bps * sms;
        const double g = fl / (ms * 1e-3) / 1e9;
        if (g > best) best = g;
      }
      printf("threads/block %4d blocks/SM %d : %.1f GFLOP/s FP64 (DFMA)\n", tpb, bps, best);
      if (best > best_all) best_all = best;
    }
  }
  printf("{\"fp64_dfma_gflops\": %.1f, \"sms\": %d}\n", best_all, sms);
  return 0;
}
