// Probe: validates the tcgen05 encodings K1 (tensor mode) relies on, on a real B200.
//   - B operand: smem, K-major, no swizzle, core matrices [kchunk][ngroup][8 rows][8 fp16]
//                (LBO = N*16 B between K chunks, SBO = 128 B between 8-row groups)
//   - A operand: (SS) smem same canonical layout, (TS) TMEM lane=row, column j = fp16x2 (k=2j,2j+1)
//   - D: TMEM fp32, lane=row, column=n; read back with tcgen05.ld.32x32b
// Prints max |err| versus a host reference for several N.  nvcc -arch=sm_100a tools/umma_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  return d;                 // layout_type 0 = no swizzle
}
__device__ __forceinline__ uint32_t make_idesc(int M, int N) {
  // c=f32 (1<<4), a=f16 (0<<7), b=f16 (0<<10), K-major both, n>>3 @17, m>>4 @24
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}\n" ::"r"(
          smem_u32(b)),
      "r"(parity));
}

// mode 0: SS, mode 1: TS
__global__ void probe(const __half* A, const __half* Bw, float* D, int N, int K, int mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  uint8_t* sB = smem;                 // N*K*2 bytes
  uint8_t* sA = smem + 65536;         // 128*K*2 bytes
  const int tid = threadIdx.x, warp = tid >> 5;
  // stage B and A in canonical layout: element (r,k) at ((k/8)*(R/8) + r/8)*128 + (r%8)*16 + (k%8)*2
  for (int i = tid; i < N * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    *(__half*)(sB + ((k / 8) * (N / 8) + r / 8) * 128 + (r % 8) * 16 + (k % 8) * 2) = Bw[i];
  }
  for (int i = tid; i < 128 * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    *(__half*)(sA + ((k / 8) * 16 + r / 8) * 128 + (r % 8) * 16 + (k % 8) * 2) = A[i];
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) mbar_init(&bar, 1);
  asm volatile("fence.proxy.async.shared::cta;");   // generic-proxy smem writes -> async proxy (UMMA reads)
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tb = tmem_base_s;
  const uint32_t tD = tb, tA = tb + 256;
  if (mode == 1 && warp < 4) {
    // each thread = one row (lane 32*warp + lane); write K/2 packed columns
    const int row = tid;
    for (int j0 = 0; j0 < K / 2; j0 += 8) {
      uint32_t v[8];
      for (int j = 0; j < 8; ++j) {
        __half2 h = __halves2half2(A[row * K + 2 * (j0 + j)], A[row * K + 2 * (j0 + j) + 1]);
        v[j] = *(uint32_t*)&h;
      }
      uint32_t addr = tA + ((uint32_t)(warp * 32) << 16) + j0;
      asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(addr), "r"(v[0]),
                   "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]));
    }
    asm volatile("tcgen05.wait::st.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  if (warp == 4 && (tid & 31) == 0) {
    const uint32_t idesc = make_idesc(128, N);
    for (int s = 0; s < K / 16; ++s) {
      uint64_t bdesc = make_desc(smem_u32(sB) + s * 2 * (N / 8) * 128, (N / 8) * 128, 128);
      uint32_t acc = s > 0;
      if (mode == 0) {
        uint64_t adesc = make_desc(smem_u32(sA) + s * 2 * 16 * 128, 16 * 128, 128);
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(tD),
                     "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc));
      } else {
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}" ::"r"(tD),
                     "r"(tA + s * 8), "l"(bdesc), "r"(idesc), "r"(acc));
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
  }
  if (warp < 4) {
    mbar_wait(&bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;");
    const int row = tid;
    for (int c0 = 0; c0 < N; c0 += 8) {
      uint32_t v[8];
      uint32_t addr = tD + ((uint32_t)(warp * 32) << 16) + c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "r"(addr));
      asm volatile("tcgen05.wait::ld.sync.aligned;");
      for (int j = 0; j < 8; ++j) D[row * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tb));
}

int main() {
  int fails = 0;
  const int Ns[] = {16, 64, 128, 256};
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 32768);
  for (int mode = 0; mode < 2; ++mode)
    for (int N : Ns) {
      const int K = 64;
      std::vector<__half> A(128 * K), B(N * K);
      std::vector<float> Af(128 * K), Bf(N * K), ref(128 * N), out(128 * N);
      srand(1234 + N);
      for (int i = 0; i < 128 * K; ++i) { float v = (rand() % 2001 - 1000) / 500.f; A[i] = __float2half(v); Af[i] = __half2float(A[i]); }
      for (int i = 0; i < N * K; ++i) { float v = (rand() % 2001 - 1000) / 500.f; B[i] = __float2half(v); Bf[i] = __half2float(B[i]); }
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
          double s = 0;
          for (int k = 0; k < K; ++k) s += (double)Af[m * K + k] * Bf[n * K + k];
          ref[m * N + n] = (float)s;
        }
      __half *dA, *dB; float* dD;
      cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dD, out.size() * 4);
      cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
      cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
      cudaMemset(dD, 0xff, out.size() * 4);
      probe<<<1, 160, 65536 + 32768>>>(dA, dB, dD, N, K, mode);
      cudaError_t e = cudaDeviceSynchronize();
      cudaMemcpy(out.data(), dD, out.size() * 4, cudaMemcpyDeviceToHost);
      double maxerr = 0;
      for (size_t i = 0; i < out.size(); ++i) { double d = fabs((double)out[i] - ref[i]); if (!(d <= maxerr)) maxerr = d; }
      bool ok = e == cudaSuccess && maxerr < 1e-3;
      printf("mode=%s N=%d K=%d: %s maxerr=%g (%s)\n", mode ? "TS" : "SS", N, K, ok ? "OK" : "FAIL", maxerr, cudaGetErrorString(e));
      if (!ok) { ++fails; printf("  sample out[0..3]=%g %g %g %g ref=%g %g %g %g\n", out[0], out[1], out[2], out[3], ref[0], ref[1], ref[2], ref[3]); }
      cudaFree(dA); cudaFree(dB); cudaFree(dD);
    }
  printf("probe %s\n", fails ? "FAILED" : "PASSED");
  return fails ? 1 : 0;
}
