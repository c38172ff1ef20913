// Microbenchmark: legacy mma.sync m16n8k8 tf32 rate on sm_100a (per SM per clock), to size the Gram fast path of K2.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void k_mma(float* out) {
  float c[8][4];
  unsigned a[4] = {threadIdx.x, threadIdx.x + 1, threadIdx.x + 2, threadIdx.x + 3}, b[2] = {threadIdx.x, 7u};
  for (int t = 0; t < 8; ++t) for (int i = 0; i < 4; ++i) c[t][i] = 0.f;
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int t = 0; t < 8; ++t)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  float s = 0; for (int t = 0; t < 8; ++t) for (int i = 0; i < 4; ++i) s += c[t][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
  int blocks = 148 * 2, threads = 256;
  float* out; cudaMalloc(&out, blocks * threads * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_mma<<<blocks, threads>>>(out); cudaDeviceSynchronize();
  cudaEventRecord(e0); k_mma<<<blocks, threads>>>(out); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double mmas = (double)blocks * (threads / 32) * ITERS * 8;
  printf("mma.sync m16n8k8 tf32: %.3f ms, %.1f MAC/clk/SM (%.1f TFLOP/s)\n", ms, mmas * 1024 / (ms * 1e-3) / 148 / 1.965e9,
         mmas * 2048 / (ms * 1e-3) / 1e12);
  return 0;
}
