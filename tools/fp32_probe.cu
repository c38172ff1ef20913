// Microbenchmark: scalar FFMA vs packed fma.rn.f32x2 / add.f32x2 issue throughput on sm_100a,
// plus MUFU (sqrt.approx, rcp.approx) rate.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/fp32_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void k_ffma(float* out, float a, float b) {
  float x[16];
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
  float s = 0; for (int i = 0; i < 16; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float* out, float a, float b) {
  unsigned long long x[16], aa, bb;
  asm("mov.b64 %0, {%1,%1};" : "=l"(aa) : "f"(a));
  asm("mov.b64 %0, {%1,%1};" : "=l"(bb) : "f"(b));
  for (int i = 0; i < 16; ++i) { float v = threadIdx.x + i; asm("mov.b64 %0, {%1,%1};" : "=l"(x[i]) : "f"(v)); }
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[i]) : "l"(aa), "l"(bb));
  float s = 0;
  for (int i = 0; i < 16; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_subfma2(float* out, float a, float b) {   // the direct-difference inner loop: d = p - q; acc += d*d
  unsigned long long acc[8], p[8], q;
  asm("mov.b64 %0, {%1,%2};" : "=l"(q) : "f"(a), "f"(b));
  for (int i = 0; i < 8; ++i) { float v = threadIdx.x + i; asm("mov.b64 %0, {%1,%1};" : "=l"(p[i]) : "f"(v)); acc[i] = 0; }
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      unsigned long long d;
      asm volatile("sub.f32x2 %0, %1, %2;" : "=l"(d) : "l"(p[i]), "l"(q));
      asm volatile("fma.rn.f32x2 %0, %1, %1, %0;" : "+l"(acc[i]) : "l"(d));
    }
  float s = 0;
  for (int i = 0; i < 8; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_subfma(float* out, float a, float b) {
  float acc[16], p[16];
  for (int i = 0; i < 16; ++i) { p[i] = threadIdx.x + i; acc[i] = 0; }
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 16; ++i) { float d = p[i] - ((i & 1) ? a : b); acc[i] = fmaf(d, d, acc[i]); }
  float s = 0; for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_mufu(float* out, float a) {
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i + a;
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) asm volatile("sqrt.approx.f32 %0, %0;" : "+f"(x[i]));
  float s = 0; for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_sqrt_ieee(float* out, float a) {
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i + a;
  for (int it = 0; it < ITERS; ++it)
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = sqrtf(x[i] + a);
  float s = 0; for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
  int sms = 148, blocks = sms * 4, threads = 512;
  float* out; cudaMalloc(&out, blocks * threads * 4);
  double lanes = (double)blocks * threads * ITERS;
  float t;
  t = timeit([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf("FFMA      : %.3f ms  %.1f GFMA/s per SM-clk=%.1f\n", t, lanes * 16 / t / 1e6, lanes * 16 / (t * 1e-3) / 148 / 1.965e9);
  t = timeit([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf("FFMA2     : %.3f ms  %.1f GFMA/s per SM-clk=%.1f\n", t, lanes * 32 / t / 1e6, lanes * 32 / (t * 1e-3) / 148 / 1.965e9);
  t = timeit([&] { k_subfma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf("SUB+FMA   : %.3f ms  diff-terms per SM-clk=%.1f\n", t, lanes * 16 / (t * 1e-3) / 148 / 1.965e9);
  t = timeit([&] { k_subfma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
  printf("SUB2+FMA2 : %.3f ms  diff-terms per SM-clk=%.1f\n", t, lanes * 16 / (t * 1e-3) / 148 / 1.965e9);
  t = timeit([&] { k_mufu<<<blocks, threads>>>(out, 1.0f); });
  printf("MUFU.SQRT : %.3f ms  per SM-clk=%.1f\n", t, lanes * 8 / (t * 1e-3) / 148 / 1.965e9);
  t = timeit([&] { k_sqrt_ieee<<<blocks, threads>>>(out, 1.0f); });
  printf("sqrtf IEEE: %.3f ms  per SM-clk=%.1f\n", t, lanes * 8 / (t * 1e-3) / 148 / 1.965e9);
  return 0;
}
