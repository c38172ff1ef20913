// Probe (B200): the TMEM / tcgen05 rates the two-tile rollout kernel is designed around.
//   1. tcgen05.ld throughput with 4 / 8 / 16 warps (is the 64 B/clk figure per SM or per lane quadrant?)
//   2. tcgen05.st throughput
//   3. tcgen05.mma (kind::f16, M=128, K=16) cycles per instruction for N = 16 / 64 / 128 / 256, A from TMEM (TS) and
//      from shared memory (SS), back-to-back issue by one thread
//   4. round trip of one "layer event": MMA -> commit -> epilogue warps (ld 64 cols, relu, cvt, st 32 cols) -> mbarrier
//      arrive -> issuer -> next MMA, for 1 and 2 interleaved tiles
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ndivplanning_b200/csrc tools/tmem_probe.cu -o tools/tmem_probe
#include <cstdio>
#include <cstdlib>
#include "tc_common.cuh"

using namespace ndp::tc;

__device__ unsigned long long g_out[64];

// ---- 1/2: ld / st throughput ------------------------------------------------------------------
template <int MODE>   // 0 = ld x32 pairs, 1 = st x16 pairs
__global__ void __launch_bounds__(512, 1) ldst_kernel(int reps, int slot) {
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc512(smem_u32(&tmem_slot));
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 64);
  uint32_t v[32], w[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) { v[i] = i; w[i] = i + 1; }
  __syncthreads();
  const long long t0 = clock64();
  uint32_t acc = 0;
  for (int r = 0; r < reps; ++r) {
    if (MODE == 0) {
      tmem_ld32(base, v);
      tmem_ld32(base + 32, w);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) acc ^= v[i] ^ w[i];
    } else {
      tmem_st16(base, v); tmem_st16(base + 16, v + 16);
      tmem_st16(base + 32, w); tmem_st16(base + 48, w + 16);
      tmem_st_wait();
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) g_out[slot] = (unsigned long long)(t1 - t0);
  if (acc == 0x12345678u) g_out[63] = acc;
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc512(tmem);
}

// ---- 3: MMA rate --------------------------------------------------------------------------------
template <int N, int TS>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int reps, int slot) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (65536 + 4096) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u;  // 1.0h
  if (warp == 0) tmem_alloc512(smem_u32(&tmem_slot));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = idesc_f16(N);
    const uint32_t b0 = smem_u32(smem), a0 = smem_u32(smem + 65536);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      // 8 K-slabs per "item", B slabs N*32 bytes apart (LBO = N*16)
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const uint64_t bd = smem_desc(b0 + s * N * 32, N * 16, 128);
        if (TS) mma_ts(tmem, tmem + 256 + s * 8, bd, idesc, (r | s) != 0);
        else mma_ss(tmem, smem_desc(a0, 2048, 128), bd, idesc, (r | s) != 0);
      }
    }
    const long long t_issue = clock64();
    mma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    const long long t1 = clock64();
    g_out[slot] = (unsigned long long)(t1 - t0);
    g_out[slot + 1] = (unsigned long long)(t_issue - t0);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc512(tmem);
}

// ---- 3b: MMA issue floor: descriptors precomputed in registers, accumulate flag constant, NI issuing warps (one
// elected thread each) streaming into disjoint D regions concurrently
template <int N, int NI>
__global__ void __launch_bounds__(128, 1) mma_issue_kernel(int reps, int slot) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bar[4];
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (65536 + 4096) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u;
  if (warp == 0) tmem_alloc512(smem_u32(&tmem_slot));
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_mbar_init(); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const long long t0 = clock64();
  if (warp < NI && (threadIdx.x & 31) == 0) {
    const uint32_t idesc = idesc_f16(N);
    const uint32_t b0 = smem_u32(smem);
    uint64_t bd[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) bd[s] = smem_desc(b0 + s * (N >= 128 ? 4096 : N * 32), N * 16, 128);
    const uint32_t D = tmem + warp * 128, A = tmem + 256 + warp * 64;
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int s = 0; s < 8; ++s)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(D), "r"(A + s * 8),
                     "l"(bd[s]), "r"(idesc)
                     : "memory");
    }
    mma_commit(smem_u32(&bar[warp]));
    mbar_wait(smem_u32(&bar[warp]), 0);
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) g_out[slot] = (unsigned long long)(t1 - t0);
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc512(tmem);
}

// ---- 4: layer-event round trip -------------------------------------------------------------------
// NT tiles (1 or 2); per event and tile: 8 MMAs (N=128, K=16, TS) into D_t, commit -> 8 epilogue warps per tile:
// ld 64 cols, relu, cvt fp16, st 32 cols in place, arrive -> issuer.  Issue order with two tiles: A, B, A, B ...
template <int NT>
__global__ void __launch_bounds__(32 + 256 * NT, 1) event_kernel(int events, int slot) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) unsigned long long bars[8];   // d_full[t], a_ready[t]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x2C002C00u;  // small fp16
  if (warp == 0) tmem_alloc512(smem_u32(&tmem_slot));
  if (tid == 0) {
    for (int t = 0; t < NT; ++t) { mbar_init(smem_u32(&bars[t]), 1); mbar_init(smem_u32(&bars[2 + t]), 8); }
    fence_mbar_init();
  }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const long long t0 = clock64();
  if (warp == 0) {
    const uint32_t idesc = idesc_f16(128);
    const uint32_t b0 = smem_u32(smem);
    for (int e = 0; e < events; ++e) {
      for (int t = 0; t < NT; ++t) {
        if (e > 0) { mbar_wait(smem_u32(&bars[2 + t]), (e - 1) & 1); fence_after_sync(); }
        if (elect_one()) {
          const uint32_t D = tmem + t * 256, A = tmem + t * 256 + 128 + (e & 1) * 0;
#pragma unroll
          for (int s = 0; s < 8; ++s)
            mma_ts(D, A + s * 8, smem_desc(b0 + s * 4096, 2048, 128), idesc, s != 0);
          mma_commit(smem_u32(&bars[t]));
        }
        __syncwarp();
      }
    }
    for (int t = 0; t < NT; ++t) mbar_wait(smem_u32(&bars[2 + t]), (events - 1) & 1);
  } else {
    const int ew = warp - 1, t = ew >> 3, q = ew & 3, h = (ew >> 2) & 1;
    const uint32_t addr = tmem + t * 256 + ((uint32_t)(q * 32) << 16) + 64 * h;
    for (int e = 0; e < events; ++e) {
      mbar_wait(smem_u32(&bars[t]), e & 1);
      fence_after_sync();
      uint32_t v[32], w[32], hi[32];
      tmem_ld32(addr, v);
      tmem_ld32(addr + 32, w);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        hi[j] = pack_hi(fmaxf(__uint_as_float(v[2 * j]), 0.f), fmaxf(__uint_as_float(v[2 * j + 1]), 0.f));
        hi[16 + j] = pack_hi(fmaxf(__uint_as_float(w[2 * j]), 0.f), fmaxf(__uint_as_float(w[2 * j + 1]), 0.f));
      }
      tmem_st16(addr + 128 - 64 * h + 32 * h, hi);          // operand region of the tile: cols [128,192)
      tmem_st16(addr + 128 - 64 * h + 32 * h + 16, hi + 16);
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[2 + t]));
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (tid == 0) g_out[slot] = (unsigned long long)(t1 - t0);
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc512(tmem);
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main() {
  unsigned long long out[64];
  const int reps = 2000;
  // 1/2
  for (int nw : {4, 8, 16}) {
    ldst_kernel<0><<<1, nw * 32>>>(reps, 0);
    ldst_kernel<1><<<1, nw * 32>>>(reps, 1);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpyFromSymbol(out, g_out, sizeof(out)));
    const double bytes = (double)nw * 32 * 64 * 4 * reps;
    printf("warps=%2d  tcgen05.ld: %7.1f B/clk (%.0f clk per 64-col warp read)   tcgen05.st: %7.1f B/clk\n", nw,
           bytes / out[0], (double)out[0] / reps, bytes / out[1]);
  }
  // 3
  const int smem3 = 65536 + 4096;
#define RATE(N, TS)                                                                                         \
  CK(cudaFuncSetAttribute(mma_rate_kernel<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem3));     \
  mma_rate_kernel<N, TS><<<1, 128, smem3>>>(200, 2);                                                        \
  CK(cudaDeviceSynchronize());                                                                              \
  CK(cudaMemcpyFromSymbol(out, g_out, sizeof(out)));                                                        \
  printf("mma N=%3d %s: %6.1f clk/MMA total, %6.1f clk/MMA issue  (floor N/2 = %d)\n", N, TS ? "TS" : "SS",  \
         (double)out[2] / 1600, (double)out[3] / 1600, N / 2);
  RATE(16, 1) RATE(64, 1) RATE(128, 1) RATE(256, 1) RATE(64, 0) RATE(128, 0)
#define ISSUE(N, NI)                                                                                        \
  CK(cudaFuncSetAttribute(mma_issue_kernel<N, NI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem3));    \
  mma_issue_kernel<N, NI><<<1, 128, smem3>>>(200, 6);                                                       \
  CK(cudaDeviceSynchronize());                                                                              \
  CK(cudaMemcpyFromSymbol(out, g_out, sizeof(out)));                                                        \
  printf("mma issue floor N=%3d, %d issuer warp(s), descriptors in registers: %6.1f clk per MMA per issuer, %6.1f clk per MMA aggregate\n", \
         N, NI, (double)out[6] / 1600, (double)out[6] / 1600 / NI);
  ISSUE(16, 1) ISSUE(64, 1) ISSUE(128, 1) ISSUE(16, 2) ISSUE(64, 2) ISSUE(128, 2)
  // 4
  CK(cudaFuncSetAttribute(event_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(event_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  event_kernel<1><<<1, 32 + 256, 65536>>>(500, 4);
  CK(cudaDeviceSynchronize());
  event_kernel<2><<<1, 32 + 512, 65536>>>(500, 5);
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(out, g_out, sizeof(out)));
  printf("layer event (8 MMAs N=128 = 512 clk of tensor work + epilogue 64->32 cols): 1 tile %.0f clk/event, "
         "2 tiles interleaved %.0f clk/event-pair (%.0f per tile-event)\n",
         (double)out[4] / 500, (double)out[5] / 500, (double)out[5] / 1000);
  return 0;
}
