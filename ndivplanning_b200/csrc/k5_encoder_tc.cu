// K5 on the tensor cores — conv2..conv5 of models/image_autoencoder.py:36-46 (3x3, stride 2, pad 1; 64 -> 128 -> 256
// -> 512 -> 1024 channels, 9.7 GFLOP each at B = 64) as implicit GEMMs on tcgen05, same exactness protocol as K1's
// f16x3: both operands are split into fp16 hi + lo (22 significand bits) and the three products hi.hi + hi.lo + lo.hi
// accumulate in fp32 in TMEM.
//
//   Y[m, n] = sum_k A(m, k) W[n, k],   m = (image, out row, out col),  k = (kernel row, kernel col, in channel)
//
// * Operand scaling (what keeps the lo halves normal numbers, as in K2): weights are scaled per OUTPUT CHANNEL by a
//   power of two (max |w| * s in [256, 512)) when the parameters are loaded; activations per LAYER by a power of two
//   taken from the layer's running maximum (max * s in [8192, 16384); the producing epilogue keeps the maximum with
//   one atomicMax per warp, a split pass turns the fp32 NHWC activation into two fp16 NHWC planes).  Powers of two:
//   the scaling itself is exact, the epilogue multiplies the accumulator by 1 / (s_a * s_w[n]).
// * Weights are packed once into MMA-ready tiles: [n tile of 128][k chunk of 32][hi | lo], each 8 KB in the
//   canonical K-major no-swizzle layout (core matrix = 8 channels-out x 16 B), so ONE 16 KB bulk copy (TMA engine)
//   fills a stage's B operand.
// * The im2col matrix is never materialised: a k chunk of 32 lies inside one kernel cell (Cin % 32 == 0), i.e. 32
//   contiguous channels of one input pixel = 64 B per plane; four consecutive producer lanes fetch the four 16-byte
//   segments of a row with zero-filling async copies (padding and rows beyond M arrive as zeros) straight into the
//   canonical layout (the lane mapping is measured, see the producer loop).
// * One CTA = 128 output pixels x 128 or 256 output channels (NT = 1 / 2 tiles from one gathered A tile; x a range of k chunks when the layer has few tiles: split-K
//   into partial slots that a finalize kernel adds in fixed order).  160 threads: warps 0-3 gather, then run the
//   epilogue (TMEM lane quadrant = warp); warp 4 issues 6 tcgen05.mma per stage (2 K-slabs x 3 products) and
//   releases the stage with tcgen05.commit.  6 stages x 32 KB (NT = 1) or 4 x 48 KB (NT = 2), one CTA per SM.
// * hi.hi goes into one TMEM accumulator, the two cross products (2^-11 of it) into another, added in the epilogue:
//   the tensor pipe truncates when it aligns an addend to the accumulator, so the small terms never pass through the
//   large accumulator (measured, tools/encoder_tc_check.py: 8.5e-6 -> 1.3e-6 max error at B = 64).
// Roofline note: a 128 x 128 x 32 stage moves 32 KB for 3 x 1.05 MFLOP = 96 FLOP/B; at the measured L2 cap of
// ~43 B/clk/SM that is ~4.1 kFLOP/clk/SM, half the dense fp16 tensor rate: the layer is L2-bandwidth bound by design.
#include <cuda_fp16.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "encoder_tc.cuh"

namespace ndp {
namespace enc {
using namespace ndp::tc;

constexpr int kCM = 128, kCN = 128, kCK = 32, kCThreads = 160;
// One CTA computes NT = 1 or 2 adjacent 128-channel tiles from ONE gathered A tile.  cp.async writes shared memory one
// wavefront per fetched 32-byte sector (ncu: 16 KB of gathered A = 512-640 shared wavefronts against 384 tensor cycles
// at NT = 1, profiles/r02d_ncu_conv_tc_*_raw.csv), so the gather, not the tensor pipe, bounds a 128-wide tile; NT = 2
// halves the gathered bytes per MMA.  lag(NT) gather groups stay in flight and stages(NT) > lag + 1, so the slot the
// producers wait for at iteration i was handed to the tensor pipe TWO iterations earlier and its MMAs run under the
// issue of the next loads.
__host__ __device__ constexpr int conv_stages(int NT) { return NT == 1 ? 6 : 4; }
__host__ __device__ constexpr int conv_lag(int NT) { return NT == 1 ? 4 : 2; }
static_assert(kCK == 32, "the gather maps 4 lanes x 16 B to one row of a k chunk");
constexpr uint32_t kCPlane = 128 * kCK * 2;               // one [128 x 32] fp16 operand plane (8 KB)
__host__ __device__ constexpr uint32_t conv_stage_bytes(int NT) { return (2 + 2 * NT) * kCPlane; }   // A hi | A lo | NT x (B hi | B lo)
__host__ __device__ constexpr uint32_t conv_smem(int NT) { return conv_stages(NT) * conv_stage_bytes(NT) + 128; }   // 192 KB: one CTA per SM

__device__ __forceinline__ void cp_async16z(uint32_t smem_dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// power of two s with v * s in [2^(e-1), 2^e) for v > 0 (1 for v == 0 or non-finite)
__device__ __forceinline__ float pow2_scale(float v, int e) {
  if (!(v > 0.f) || !(v < 3e38f)) return 1.f;
  int ex;
  frexpf(v, &ex);                                          // v = f * 2^ex, f in [0.5, 1)
  int k = e - ex;
  k = k > 120 ? 120 : (k < -120 ? -120 : k);
  return ldexpf(1.f, k);
}

// ---------------------------------------------------------------- weight packing (ndp_encoder_load)
// W [Cout][K] fp32 (k = (kh*KW + kw)*Cin + ci) -> wpack [Cout/128][K/32][hi 8 KB | lo 8 KB], wsinv[n] = 1 / s_w[n]
__global__ void conv_tc_wscale_kernel(const float* __restrict__ W, int K, float* __restrict__ wsinv, float* __restrict__ ws) {
  const int n = blockIdx.x;
  float mx = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) mx = fmaxf(mx, fabsf(__ldg(W + (size_t)n * K + k)));
  __shared__ float sm[32];
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmaxf(mx, sm[w]);
    const float s = pow2_scale(mx, 9);                     // max |w| * s in [256, 512)
    ws[n] = s;
    wsinv[n] = 1.f / s;
  }
}
__global__ void conv_tc_wpack_kernel(const float* __restrict__ W, int Cout, int K, const float* __restrict__ ws,
                                     uint8_t* __restrict__ wpack) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;        // one pair of consecutive k per thread
  if (e >= (size_t)Cout * K / 2) return;
  const int n = (int)(e / (K / 2)), k = (int)(e % (K / 2)) * 2;
  const float s = ws[n];
  const float w0 = __ldg(W + (size_t)n * K + k) * s, w1 = __ldg(W + (size_t)n * K + k + 1) * s;
  const uint32_t hi = pack_hi(w0, w1), lo = pack_lo(w0, w1, hi);
  const int nt = n / kCN, nn = n % kCN, kc = k / kCK, kk = k % kCK, nchunks = K / kCK;
  uint8_t* tile = wpack + ((size_t)nt * nchunks + kc) * (2 * kCPlane);
  const uint32_t off = (uint32_t)(kk / 8) * 2048 + (uint32_t)(nn / 8) * 128 + (uint32_t)(nn % 8) * 16 + (uint32_t)(kk % 8) * 2;
  *reinterpret_cast<uint32_t*>(tile + off) = hi;
  *reinterpret_cast<uint32_t*>(tile + kCPlane + off) = lo;
}

size_t conv_tc_pack_bytes(int Cout, int K) { return (size_t)Cout * K * 4; }

int conv_tc_pack(const float* W, int Cout, int K, uint8_t* wpack, float* wsinv, float* ws, cudaStream_t s) {
  NDP_CHECK_ARG(Cout % kCN == 0 && K % kCK == 0, "conv_tc_pack: Cout %% 128 and K %% 32 must be 0");
  conv_tc_wscale_kernel<<<Cout, 256, 0, s>>>(W, K, wsinv, ws);
  NDP_LAUNCH_CHECK();
  const size_t pairs = (size_t)Cout * K / 2;
  conv_tc_wpack_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, s>>>(W, Cout, K, ws, wpack);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---------------------------------------------------------------- activation split pass
// act fp32 NHWC -> hi / lo fp16 planes scaled by the layer's power of two; ainv[0] = 1 / s for the conv epilogue
__global__ void conv_tc_split_kernel(const float4* __restrict__ act, size_t n4, const float* __restrict__ amax,
                                     float* __restrict__ ainv, uint2* __restrict__ hi, uint2* __restrict__ lo) {
  const float s = pow2_scale(__ldg(amax), 14);             // max * s in [8192, 16384): far from fp16 overflow
  if (blockIdx.x == 0 && threadIdx.x == 0) ainv[0] = 1.f / s;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(act + i);
    uint2 h, l;
    h.x = pack_hi(v.x * s, v.y * s); l.x = pack_lo(v.x * s, v.y * s, h.x);
    h.y = pack_hi(v.z * s, v.w * s); l.y = pack_lo(v.z * s, v.w * s, h.y);
    hi[i] = h;
    lo[i] = l;
  }
}

int conv_tc_split(const float* act, size_t n, const float* amax, float* ainv, __half* hi, __half* lo, cudaStream_t s) {
  NDP_CHECK_ARG(n % 4 == 0, "conv_tc_split: element count must be a multiple of 4");
  const size_t n4 = n / 4;
  size_t blocks = (n4 + 255) / 256;
  const size_t cap = (size_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  conv_tc_split_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const float4*>(act), n4, amax, ainv,
                                                        reinterpret_cast<uint2*>(hi), reinterpret_cast<uint2*>(lo));
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---------------------------------------------------------------- the layer
template <int NT>
__global__ void __launch_bounds__(kCThreads, 1) conv_tc_kernel(const ConvTcP p) {
  constexpr int kCStages = conv_stages(NT), kCLag = conv_lag(NT);
  constexpr uint32_t kCStageBytes = conv_stage_bytes(NT);
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bars = sbase + kCStages * kCStageBytes;        // full[S] @0, empty[S] @8S, acc @16S, tmem slot @112
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kCStages * kCStageBytes + 112);
  constexpr uint32_t kEmpty = 8 * kCStages, kAcc = 16 * kCStages;
  const int m0 = blockIdx.x * kCM, nt0 = blockIdx.y * NT;       // first of this CTA's NT output-channel tiles
  const int c_begin = blockIdx.z * p.chunks_per_split;
  const int c_end = min(p.n_chunks, c_begin + p.chunks_per_split);
  const int nloc = c_end - c_begin;

  if (tid == 0) {
    for (int s = 0; s < kCStages; ++s) { mbar_init(bars + 8 * s, 129); mbar_init(bars + kEmpty + 8 * s, 1); }
    mbar_init(bars + kAcc, 1);
    fence_mbar_init();
  }
  // TMEM: tile j's hi.hi accumulator at columns [128 j, 128 j + 128), its cross products at [128 (NT + j), ...)
  if (warp == 4) { if (NT == 1) tmem_alloc256(smem_u32((const void*)tmem_slot)); else tmem_alloc512(smem_u32((const void*)tmem_slot)); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 4) {
    // ===================== gather producers =====================
    // Lane l of warp w fetches 16-byte segment l % 4 of rows q*32 + w*8 + l/4 (q = 0..3): the 4 lanes of a row read its
    // 64 contiguous bytes.  Measured (B = 64, 3-stage build): a row per lane 300 us for the four layers, this mapping
    // 249 us, 8 rows x one segment per quarter warp 272 us — the shared-memory side of LDGSTS takes one wavefront
    // per fetched sector, so fewer sectors per quarter warp win even though the 4 segments of a row (2048 B apart)
    // share banks.
    const int seg = lane & 3;
    int ih0[4], iw0[4];
    size_t pix0[4];
    bool rowok[4];
    uint32_t row_off[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = q * 32 + warp * 8 + (lane >> 2), mm = m0 + r;
      rowok[q] = mm < p.M;
      ih0[q] = iw0[q] = 0; pix0[q] = 0;
      if (rowok[q]) {
        const int b = mm / (p.OH * p.OW), qq = mm % (p.OH * p.OW);
        ih0[q] = (qq / p.OW) * p.stride - p.pad; iw0[q] = (qq % p.OW) * p.stride - p.pad;
        pix0[q] = (size_t)b * p.H * p.W;
      }
      row_off[q] = (uint32_t)seg * 2048 + (uint32_t)(r >> 3) * 128 + (uint32_t)(r & 7) * 16;
    }
    const uint8_t* wsrc = p.wpack + (size_t)nt0 * p.n_chunks * (2 * kCPlane);
    for (int i = 0; i < nloc; ++i) {
      const int c = c_begin + i, stage = i % kCStages;
      if (i >= kCStages) mbar_wait(bars + kEmpty + 8 * stage, (uint32_t)((i / kCStages) - 1) & 1u);
      const int k0 = c * kCK, cell = k0 / p.Cin, ci0 = k0 - cell * p.Cin + seg * 8;
      const int kh = cell / p.KW, kw = cell - kh * p.KW;
      const uint32_t a_hi = sbase + stage * kCStageBytes, a_lo = a_hi + kCPlane;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int ih = ih0[q] + kh, iw = iw0[q] + kw;
        const bool ok = rowok[q] && ih >= 0 && ih < p.H && iw >= 0 && iw < p.W;
        const size_t off = ok ? ((pix0[q] + (size_t)ih * p.W + iw) * p.Cin + ci0) : 0;
        cp_async16z(a_hi + row_off[q], p.in_hi + off, ok);
        cp_async16z(a_lo + row_off[q], p.in_lo + off, ok);
      }
      if (tid == 0) {
        mbar_expect_tx(bars + 8 * stage, NT * 2 * kCPlane);
#pragma unroll
        for (int j = 0; j < NT; ++j)
          bulk_g2s(sbase + stage * kCStageBytes + (2 + 2 * j) * kCPlane,
                   wsrc + ((size_t)j * p.n_chunks + c) * (2 * kCPlane), 2 * kCPlane, bars + 8 * stage);
      }
      cp_commit();
      if (i >= kCLag) {                                     // the group of chunk i - kCLag has landed
        cp_wait<kCLag>();
        fence_proxy_async();
        mbar_arrive(bars + 8 * ((i - kCLag) % kCStages));
      }
    }
    cp_wait<0>();                                           // tail: the last kCLag groups
    fence_proxy_async();
    for (int d = nloc - kCLag > 0 ? nloc - kCLag : 0; d < nloc; ++d) mbar_arrive(bars + 8 * (d % kCStages));
    // ===================== epilogue: row r of the tile, 128 output channels =====================
    const int m = m0 + tid;                                 // epilogue: thread = TMEM lane = row of the tile
    const bool rvalid = m < p.M;
    mbar_wait(bars + kAcc, 0);
    fence_after_sync();
    const float ainv = __ldg(p.ainv);
    float* dst = (p.splits > 1 ? p.part + (size_t)blockIdx.z * p.M * p.Cout : p.out) + (size_t)m * p.Cout;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    float mx = 0.f;
#pragma unroll 1
    for (int cc = 0; cc < 4 * NT; ++cc) {                   // 32 output channels at a time
      const int j = cc >> 2, n0 = (nt0 + j) * kCN + (cc & 3) * 32;
      uint32_t v[32], vc[32];
      tmem_ld32(lane_base + j * 128 + (cc & 3) * 32, v);
      tmem_ld32(lane_base + (NT + j) * 128 + (cc & 3) * 32, vc);
      tmem_ld_wait();
      if (rvalid) {
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4) {
          float o[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int n = n0 + q4 * 4 + u;
            float x = (__uint_as_float(v[q4 * 4 + u]) + __uint_as_float(vc[q4 * 4 + u])) * (ainv * __ldg(p.wsinv + n));
            if (p.splits == 1) { x += __ldg(p.bias + n); if (p.relu) x = fmaxf(x, 0.f); mx = fmaxf(mx, x); }
            o[u] = x;
          }
          *reinterpret_cast<float4*>(dst + n0 + q4 * 4) = make_float4(o[0], o[1], o[2], o[3]);
        }
      }
    }
    if (p.splits == 1 && p.amax_out) {
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (lane == 0) atomicMax(reinterpret_cast<int*>(p.amax_out), __float_as_int(mx));    // values >= 0
    }
  } else if (elect_one()) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = idesc_f16(kCN);
    for (int i = 0; i < nloc; ++i) {
      const int stage = i % kCStages;
      mbar_wait(bars + 8 * stage, (uint32_t)(i / kCStages) & 1u);
      fence_after_sync();
      const uint32_t a_hi = sbase + stage * kCStageBytes, a_lo = a_hi + kCPlane;
      // hi.hi into one accumulator, the two cross products (2^-11 of it) into another: the tensor pipe truncates when it
      // aligns an addend to the accumulator, so the main chain is kept to K/16 additions and the small terms never
      // pass through the large accumulator (measured: tools/encoder_tc_check.py)
#pragma unroll
      for (int s = 0; s < kCK / 16; ++s) {
        const uint64_t ah = smem_desc(a_hi + s * 4096, 2048, 128), al = smem_desc(a_lo + s * 4096, 2048, 128);
#pragma unroll
        for (int j = 0; j < NT; ++j) {
          const uint32_t b_hi = a_lo + (1 + 2 * j) * kCPlane, b_lo = b_hi + kCPlane;
          const uint64_t bh = smem_desc(b_hi + s * 4096, 2048, 128), bl = smem_desc(b_lo + s * 4096, 2048, 128);
          mma_ss(tmem + j * 128, ah, bh, idesc, (i | s) != 0);
          mma_ss(tmem + (NT + j) * 128, ah, bl, idesc, (i | s) != 0);
          mma_ss(tmem + (NT + j) * 128, al, bh, idesc, 1);
        }
      }
      mma_commit(bars + kEmpty + 8 * stage);                 // the stage may be refilled
    }
    mma_commit(bars + kAcc);                                 // accumulators complete
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 4) { if (NT == 1) tmem_dealloc256(tmem); else tmem_dealloc512(tmem); }
}

// out[m][n] = act(sum_s part[s][m][n] + bias[n]), splits added in order; keeps the running maximum
__global__ void __launch_bounds__(256) conv_tc_finalize_kernel(const float4* __restrict__ part, int splits, size_t MN4, int N,
                                                               const float* __restrict__ bias, int relu,
                                                               float4* __restrict__ out, float* __restrict__ amax) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float mx = 0.f;
  if (i < MN4) {
    float4 a = __ldg(part + i);
    for (int s = 1; s < splits; ++s) {
      const float4 b = __ldg(part + (size_t)s * MN4 + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    const int n = (int)((i * 4) % N);
    a.x += __ldg(bias + n); a.y += __ldg(bias + n + 1); a.z += __ldg(bias + n + 2); a.w += __ldg(bias + n + 3);
    if (relu) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
    out[i] = a;
    mx = fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w));
  }
  if (amax) {
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(reinterpret_cast<int*>(amax), __float_as_int(mx));
  }
}

int conv_tc_forward(ConvTcP p, size_t part_floats, cudaStream_t s) {
  NDP_CHECK_ARG(p.Cin % kCK == 0 && p.Cout % kCN == 0 && p.K % kCK == 0, "conv_tc: Cin %% 32, Cout %% 128");
  p.n_chunks = p.K / kCK;
  const int mt = (p.M + kCM - 1) / kCM;
  static const bool nt1 = getenv("NDP_ENC_NT1") != nullptr;               // diagnostics: 128-wide tiles only
  const int NT = (!nt1 && p.Cout % (2 * kCN) == 0) ? 2 : 1;
  const int ntl = p.Cout / (kCN * NT), tiles = mt * ntl;
  int splits = 1;
  if (tiles < num_sms()) {                                 // one resident CTA per SM
    splits = num_sms() / tiles;
    const int max_by_k = p.n_chunks / 12 > 0 ? p.n_chunks / 12 : 1;        // at least 12 chunks (K = 384) per split
    if (splits > max_by_k) splits = max_by_k;
    while (splits > 1 && (size_t)splits * p.M * p.Cout > part_floats) --splits;
    if (splits < 1) splits = 1;
  }
  p.chunks_per_split = (p.n_chunks + splits - 1) / splits;
  p.splits = (p.n_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  static std::atomic<unsigned long long> attr_devs{0};
  if (first_use_on_device(attr_devs)) {
    NDP_CUDA(cudaFuncSetAttribute(conv_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)conv_smem(1)));
    NDP_CUDA(cudaFuncSetAttribute(conv_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)conv_smem(2)));
  }
  if (NT == 1) conv_tc_kernel<1><<<dim3(mt, ntl, p.splits), kCThreads, conv_smem(1), s>>>(p);
  else conv_tc_kernel<2><<<dim3(mt, ntl, p.splits), kCThreads, conv_smem(2), s>>>(p);
  NDP_LAUNCH_CHECK();
  if (p.splits > 1) {
    const size_t MN4 = (size_t)p.M * p.Cout / 4;
    conv_tc_finalize_kernel<<<(unsigned)((MN4 + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(p.part), p.splits, MN4,
                                                                         p.Cout, p.bias, p.relu,
                                                                         reinterpret_cast<float4*>(p.out), p.amax_out);
    NDP_LAUNCH_CHECK();
  }
  return NDP_OK;
}

}  // namespace enc
}  // namespace ndp
