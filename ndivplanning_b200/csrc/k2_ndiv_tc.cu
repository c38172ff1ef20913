// K2 fast path — pairwise distances in Gram form on the 5th-gen tensor cores (north_star:
// ||a||^2 + ||b||^2 - 2ab with the normalisation and hinge epilogue fused), reference diversity.py:8-41.
//
//   d2_ij = n_i + n_j - 2 <c_i, c_j>,   c = v - mean_group(v)   (centering keeps n comparable to d2),
//   <c_i,c_j> by tcgen05.mma kind::f16 on fp16 split operands c = hi + lo (22 significand bits):
//   [hi|hi|lo] x [hi|lo|hi]^T, fp32 accumulation in TMEM (one 128x128 tile per matrix and j-tile).
//
// The Gram form loses digits when d2 << n_i + n_j (near-duplicate outputs — the regime the loss exists
// for), so this path only PROPOSES a result: every pair is checked against d2 >= 2^-6 (n_i + n_j); if any
// valid off-diagonal pair fails, a device flag is raised and the exact direct-difference kernels of
// k2_ndiv.cu (launched right behind, all of their blocks exit at once when the flag is clear) recompute the
// call.  For pairs that pass, the relative error of d is <= 2^-21 * 2^6 / 2 = 1.5e-5 (at the threshold;
// 5e-7 typical), inside north_star's 1e-4.  The diagonal is exactly 0 by construction.
//
// Layout: a pre-pack kernel writes, per group and 128-row tile, the MMA-ready shared-memory images
// (K-major, no swizzle, 8x16B core matrices [k-chunk][row-group]) of the A operand [hi|hi|lo] and the
// B operand [hi|lo|hi] for x and z plus the centred norms {nx, nz}; the main kernel moves them with 1-D
// bulk copies (TMA engine), so no thread ever touches operand bytes.  One CTA = 128 rows x a range of
// 128-column tiles; 8 epilogue warps (TMEM lane quadrant = warp % 4, column half = warp / 4), two CTAs
// per SM (256 TMEM columns each) overlap one another's copy / MMA / epilogue phases.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"

namespace ndp {
using namespace tc;

static constexpr int kGT = 128;                 // rows / columns per tile
static constexpr int kGNT = 256;                // threads per CTA
static constexpr float kIllTau = 1.f / 64.f;    // conditioning threshold d2 >= tau * (n_i + n_j)

struct GramDims {
  int sx, sz;          // K=16 slabs of the x / z operands (3 * padded dims / 16)
  int img_bytes;       // bytes of one operand image (x slabs then z slabs), 4096 per slab
};
__host__ __device__ inline GramDims gram_dims(int Dx, int Dz) {
  GramDims g;
  g.sx = 3 * ((Dx + 15) / 16);
  g.sz = 3 * ((Dz + 15) / 16);
  g.img_bytes = (g.sx + g.sz) * 4096;
  return g;
}

static constexpr int kXTBytes = 16384;          // per tile: transposed x operand of the gradient product, hi + lo
static constexpr int kGNS = 130;                // float2 entries per tile in the norms array: 128 rows {nx, nz},
                                                // entry 128 = the tile's operand scales {sx, sz}, entry 129 = pad
__device__ __forceinline__ float pick_scale(float m) {   // 2^(9 - e) with m = f * 2^e, f in [0.5, 1); degenerate data -> 1
  if (!(m > 0.f) || !(m < 3e38f)) return 1.f;
  int e;
  frexpf(m, &e);
  int sh = 9 - e;
  sh = sh > 60 ? 60 : (sh < -60 ? -60 : sh);
  return ldexpf(1.f, sh);
}

// byte offset of element (row r, K index kk) inside an operand image whose matrix starts at slab0
__device__ __forceinline__ uint32_t img_off(int slab0, int r, int kk) {
  return (uint32_t)((slab0 + (kk >> 4)) * 4096 + ((kk >> 3) & 1) * 2048 + (r >> 3) * 128 + (r & 7) * 16 + (kk & 7) * 2);
}

// grid (tiles, N), 256 threads.  imgA / imgB: [N][tiles][img_bytes]; norms: [N][tiles][kGNS] float2.
// Centre = mean of the group's first min(k,128) rows (any fixed vector is exact — distances are translation
// invariant — it only has to be central enough to keep ||c||^2 comparable to d^2); every block recomputes it in
// the same order, so all tiles agree bit for bit.  Operand scale: a power of two PER TILE (max |c| * s in
// [256, 512)), so the fp16 "lo" halves stay normal numbers whatever the magnitude of the data; the main kernel
// divides the accumulator by s_I * s_J.
__global__ void __launch_bounds__(256)
ndiv_prepack_kernel(const float* __restrict__ x, const float* __restrict__ z, int k, int Dx, int Dz,
                    uint8_t* __restrict__ imgA, uint8_t* __restrict__ imgB, uint8_t* __restrict__ imgT,
                    float* __restrict__ mu_out, float2* __restrict__ norms, int* __restrict__ ill) {
  const int n = blockIdx.y, tile = blockIdx.x, tiles = gridDim.x;
  if (n == 0 && tile == 0 && blockIdx.z == 0 && threadIdx.x == 0) *ill = 0;   // the main kernels raise it (stream order)
  const GramDims gd = gram_dims(Dx, Dz);
  uint8_t* A = imgA + ((size_t)n * tiles + tile) * gd.img_bytes;
  uint8_t* B = imgB + ((size_t)n * tiles + tile) * gd.img_bytes;
  __shared__ float mu[64];
  __shared__ float s_part[4][64];
  __shared__ float s_mx[256], s_mz[256];
  __shared__ float s_sc[2];
  {
    const int d = threadIdx.x & 63, rq = threadIdx.x >> 6, D = Dx + Dz;
    const int ns = k < kGT ? k : kGT;
    float acc = 0.f;
    if (d < D) {
      const float* v = d < Dx ? x + (size_t)n * k * Dx + d : z + (size_t)n * k * Dz + (d - Dx);
      const int st = d < Dx ? Dx : Dz;
      for (int r = rq; r < ns; r += 4) acc += __ldg(v + (size_t)r * st);
    }
    s_part[rq][d] = acc;
    __syncthreads();
    if (threadIdx.x < 64) mu[threadIdx.x] = ((s_part[0][d] + s_part[1][d]) + (s_part[2][d] + s_part[3][d])) / (float)ns;
    __syncthreads();
  }
  {
    const int r0 = tile * kGT, r1 = (r0 + kGT) < k ? (r0 + kGT) : k;
    float mx = 0.f, mz = 0.f;
    for (int i = threadIdx.x; i < (r1 - r0) * Dx; i += 256)
      mx = fmaxf(mx, fabsf(__ldg(x + ((size_t)n * k + r0) * Dx + i) - mu[i % Dx]));
    for (int i = threadIdx.x; i < (r1 - r0) * Dz; i += 256)
      mz = fmaxf(mz, fabsf(__ldg(z + ((size_t)n * k + r0) * Dz + i) - mu[Dx + i % Dz]));
    s_mx[threadIdx.x] = mx; s_mz[threadIdx.x] = mz;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
        s_mx[threadIdx.x] = fmaxf(s_mx[threadIdx.x], s_mx[threadIdx.x + o]);
        s_mz[threadIdx.x] = fmaxf(s_mz[threadIdx.x], s_mz[threadIdx.x + o]);
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      s_sc[0] = pick_scale(s_mx[0]); s_sc[1] = pick_scale(s_mz[0]);
      if (blockIdx.z == 0) {
        norms[((size_t)n * tiles + tile) * kGNS + kGT] = make_float2(s_sc[0], s_sc[1]);
        norms[((size_t)n * tiles + tile) * kGNS + kGT + 1] = make_float2(0.f, 0.f);
      }
    }
    __syncthreads();
  }
  const int gx = (Dx + 15) / 16 * 2, gz = (Dz + 15) / 16 * 2;       // 8-dim groups per row (padded)
  const bool role_t = blockIdx.z == 1;                // second block of the tile: transposed image only
  for (int it = threadIdx.x; !role_t && it < kGT * (gx + gz); it += blockDim.x) {
    const int r = it % kGT, g = it / kGT;
    const bool isx = g < gx;
    const int g8 = isx ? g : g - gx;
    const int D = isx ? Dx : Dz, Dp = isx ? gx * 8 : gz * 8;
    const int row = tile * kGT + r;
    const float* src = isx ? x + ((size_t)n * k + row) * Dx : z + ((size_t)n * k + row) * Dz;
    const float* m = isx ? mu : mu + Dx;
    const float sc = isx ? s_sc[0] : s_sc[1];
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int d0 = g8 * 8 + 2 * e;
      const float c0 = (row < k && d0 < D) ? (__ldg(src + d0) - m[d0]) * sc : 0.f;
      const float c1 = (row < k && d0 + 1 < D) ? (__ldg(src + d0 + 1) - m[d0 + 1]) * sc : 0.f;
      hi[e] = pack_hi(c0, c1);
      lo[e] = pack_lo(c0, c1, hi[e]);
    }
    const int slab0 = isx ? 0 : gd.sx;
    const uint4 H = make_uint4(hi[0], hi[1], hi[2], hi[3]), L = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    const int kk = g8 * 8;
    // A operand K order [hi | hi | lo], B operand [hi | lo | hi]
    *reinterpret_cast<uint4*>(A + img_off(slab0, r, kk)) = H;
    *reinterpret_cast<uint4*>(A + img_off(slab0, r, Dp + kk)) = H;
    *reinterpret_cast<uint4*>(A + img_off(slab0, r, 2 * Dp + kk)) = L;
    *reinterpret_cast<uint4*>(B + img_off(slab0, r, kk)) = H;
    *reinterpret_cast<uint4*>(B + img_off(slab0, r, Dp + kk)) = L;
    *reinterpret_cast<uint4*>(B + img_off(slab0, r, 2 * Dp + kk)) = H;
  }
  // B operand of the gradient product  G[i, d] = sum_j What_ij c_j[d]:  rows = d (N = 32), K = j (128), K-major core
  // matrices [j/8][d/8][8 d][8 j]; fp16 hi (8 KB) then lo (8 KB) of the same scaled centred values as above.
  if (role_t) {
    uint8_t* T = imgT + ((size_t)n * tiles + tile) * kXTBytes;
    for (int it = threadIdx.x; it < 16 * 32; it += blockDim.x) {
      const int d = it & 31, jg = it >> 5;
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int r0 = tile * kGT + jg * 8 + 2 * e;
        const float c0 = (r0 < k && d < Dx) ? (__ldg(x + ((size_t)n * k + r0) * Dx + d) - mu[d]) * s_sc[0] : 0.f;
        const float c1 = (r0 + 1 < k && d < Dx) ? (__ldg(x + ((size_t)n * k + r0 + 1) * Dx + d) - mu[d]) * s_sc[0] : 0.f;
        hi[e] = pack_hi(c0, c1);
        lo[e] = pack_lo(c0, c1, hi[e]);
      }
      const uint32_t off = (uint32_t)((jg * 4 + (d >> 3)) * 128 + (d & 7) * 16);
      *reinterpret_cast<uint4*>(T + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(T + 8192 + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    if (tile == 0 && threadIdx.x < 64) mu_out[(size_t)n * 64 + threadIdx.x] = threadIdx.x < Dx ? mu[threadIdx.x] : 0.f;
  }
  if (!role_t && threadIdx.x < kGT) {                // centred squared norms, fixed summation order
    const int row = tile * kGT + threadIdx.x;
    float nx = 0.f, nz = 0.f;
    if (row < k) {
      const float* xr = x + ((size_t)n * k + row) * Dx;
      const float* zr = z + ((size_t)n * k + row) * Dz;
      for (int d = 0; d < Dx; ++d) { const float c = __ldg(xr + d) - mu[d]; nx = fmaf(c, c, nx); }
      for (int d = 0; d < Dz; ++d) { const float c = __ldg(zr + d) - mu[Dx + d]; nz = fmaf(c, c, nz); }
    }
    norms[((size_t)n * tiles + tile) * kGNS + threadIdx.x] = make_float2(nx, nz);
  }
}

// MUFU.SQRT (<= 1 ulp-class error, same order as the Gram form's own rounding); IEEE sqrtf runs at 9.7 /clk/SM against
// 15.8 for the approximation (tools/fp32_probe.cu) and the epilogue is MUFU-bound.  The exact kernels keep sqrtf.
__device__ __forceinline__ float sqrt_approx(float v) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }

typedef unsigned long long u64;   // packed fp32 pair (FFMA2)
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ u64 dup2(float v) { u64 r; asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(v)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }

struct GramParams {
  const uint8_t* imgA;
  const uint8_t* imgB;
  const float2* norms;
  const uint8_t* imgT;         // transposed x operand images (gradient product, pass 2)
  const float* mu;             // [N][64] centre of the x features
  const float* x;              // original outputs (pass 2)
  const float* sx;             // row sums (pass 2)
  const float* sz;
  int k, Dx, Dz, tiles, tps;
  int row0, rows;              // rows owned by this call (row0 is a multiple of 128)
  float margin;
  float* psx;                  // pass 1: partial row sums [JS][N][k]
  float* psz;
  float* loss_part;            // pass 2: [N][JS][row tiles of the call]
  float* gpart;                // pass 2: [JS][N][k][Dx] (nullable: loss only)
  int* ill;                    // raised when a pair fails the conditioning test
};

struct GramAcc {
  float rsx, rsz, lsum, sumW, illm;
  float g[16];                                         // gradient accumulators: dims hf*16 .. hf*16+15 of this row
};

// one pair (i, j): distances from the two Gram accumulators, then row sums (pass 1) or hinge / What / gradient
// (pass 2).  EDGE: the tile contains the diagonal, padded columns or rows outside the caller's block.
template <bool PASS2, bool EDGE, bool CHECK>
__device__ __forceinline__ void gram_pair(GramAcc& a, float accx, float accz, float2 ni, float2 nj, float cx, float cz,
                                          float2 ri, float2 rj, float margin, bool valid, bool ivalid, bool off,
                                          bool grad, float alpha, float& ws) {
  const float sxn = ni.x + nj.x, szn = ni.y + nj.y;
  float d2x = fmaf(cx, accx, sxn);
  float d2z = fmaf(cz, accz, szn);
  if (CHECK) {                                         // conditioning: min over pairs of d2 - tau (n_i + n_j)
    const float m = fminf(fmaf(-kIllTau, sxn, d2x), fmaf(-kIllTau, szn, d2z));
    a.illm = fminf(a.illm, (!EDGE || (valid && ivalid && off)) ? m : 0.f);
  }
  d2x = fmaxf(d2x, 0.f);
  d2z = fmaxf(d2z, 0.f);
  if (EDGE && !off) { d2x = 0.f; d2z = 0.f; }          // the diagonal is exactly 0
  const float ddx = sqrt_approx(d2x), ddz = sqrt_approx(d2z);
  if (!PASS2) {
    if (!EDGE || valid) { a.rsx += ddx; a.rsz += ddz; }
  } else {
    const float hij = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, ri.y)), __fmul_rn(ddx, ri.x));
    const float hji = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, rj.y)), __fmul_rn(ddx, rj.x));
    const bool ok = !EDGE || (valid && ivalid);
    if (ok) a.lsum += !(hij <= 0.f) ? hij : 0.f;
    if (grad) {
      const float wij = (hij > 0.f ? ri.x : 0.f) + (hji > 0.f ? rj.x : 0.f);
      const float w = (ok && ddx > 0.f) ? __fdividef(-wij, ddx) : 0.f;
      a.sumW += w;
      ws = w * alpha;                                  // row-scaled to O(1) for the fp16 A operand of the product
      a.illm = fminf(a.illm, 60000.f - fabsf(ws));     // would not fit fp16: let the exact kernels take the call
    }
  }
}

// grid (row tiles of the call, JS, N), 256 threads, 2 CTAs / SM
template <bool PASS2, bool CHECK>
__global__ void __launch_bounds__(kGNT, 2) ndiv_gram_kernel(const GramParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const GramDims gd = gram_dims(p.Dx, p.Dz);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = blockIdx.z, js = blockIdx.y, N = gridDim.z;
  const int itile = p.row0 / kGT + blockIdx.x;
  const int iend = (p.row0 + p.rows) < p.k ? (p.row0 + p.rows) : p.k;
  const int jt0 = js * p.tps, jt1 = (jt0 + p.tps) < p.tiles ? (jt0 + p.tps) : p.tiles;
  const bool grad = PASS2 && p.gpart != nullptr;

  uint8_t* sA = smem;
  uint8_t* sB = smem + gd.img_bytes;
  uint8_t* sXT = smem + 2 * gd.img_bytes;                                    // transposed x_j operand (pass 2)
  const uint32_t xj_bytes = PASS2 ? (uint32_t)kXTBytes : 0u;
  float2* snj = reinterpret_cast<float2*>(smem + 2 * gd.img_bytes + xj_bytes);   // {nx_j, nz_j}
  float2* srj = snj + kGNS;                                                  // {1/sx_j, 1/sz_j}
  float2* sred = srj + kGT;                                                  // [2 halves][128 rows]
  float* slos = reinterpret_cast<float*>(sred + 2 * kGT);                    // [8 warps]
  uint8_t* sbar = reinterpret_cast<uint8_t*>(slos + 8);
  const uint32_t bar_a = smem_u32(sbar), bar_full = bar_a + 8, bar_d = bar_a + 16, bar_g = bar_a + 24;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sbar + 32);

  if (tid == 0) {
    mbar_init(bar_a, 1); mbar_init(bar_full, 1); mbar_init(bar_d, 1); mbar_init(bar_g, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc256(smem_u32((const void*)tmem_slot));
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  const size_t img_i = ((size_t)n * p.tiles + itile) * gd.img_bytes;
  if (tid == 0) {
    mbar_expect_tx(bar_a, (uint32_t)gd.img_bytes);
    bulk_g2s(smem_u32(sA), p.imgA + img_i, (uint32_t)gd.img_bytes, bar_a);
  }

  // epilogue role: TMEM lane quadrant q, column half hf; this thread owns row `row` of the tile
  const int q = warp & 3, hf = warp >> 2, row = q * 32 + lane;
  const int i = itile * kGT + row;
  const bool ivalid = i >= p.row0 && i < iend;
  const float2 ni = __ldg(p.norms + ((size_t)n * p.tiles + itile) * kGNS + row);
  const float2 sI = __ldg(p.norms + ((size_t)n * p.tiles + itile) * kGNS + kGT);   // operand scales of the i tile
  float2 ri = make_float2(0.f, 0.f);
  if (PASS2 && i < p.k) ri = make_float2(1.f / __ldg(p.sx + (size_t)n * p.k + i), 1.f / __ldg(p.sz + (size_t)n * p.k + i));
  float alpha = 1.f;                                   // power of two ~ sx_i^2 / k: What * alpha = O(mean d / d)
  if (PASS2 && i < p.k) {
    const float sxi = __ldg(p.sx + (size_t)n * p.k + i);
    const float t = sxi * sxi / (float)p.k;
    if (t > 0.f && t < 3e38f) { int e; frexpf(t, &e); e = e > 100 ? 100 : (e < -100 ? -100 : e); alpha = ldexpf(1.f, e - 1); }
  }
  const uint32_t idesc = idesc_f16(128);
  const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);

  GramAcc acc;
  acc.rsx = acc.rsz = acc.lsum = acc.sumW = acc.illm = 0.f;
#pragma unroll
  for (int d = 0; d < 16; ++d) acc.g[d] = 0.f;
  const bool rows_full = itile * kGT >= p.row0 && itile * kGT + kGT <= iend;
  uint32_t ph = 0;

  for (int jt = jt0; jt < jt1; ++jt) {
    const int j0 = jt * kGT;
    if (tid == 0) {
      const uint32_t xb = grad ? (uint32_t)kXTBytes : 0u;
      mbar_expect_tx(bar_full, (uint32_t)gd.img_bytes + kGNS * 8 + xb);
      bulk_g2s(smem_u32(sB), p.imgB + ((size_t)n * p.tiles + jt) * gd.img_bytes, (uint32_t)gd.img_bytes, bar_full);
      bulk_g2s(smem_u32(snj), p.norms + ((size_t)n * p.tiles + jt) * kGNS, kGNS * 8, bar_full);
      if (grad) bulk_g2s(smem_u32(sXT), p.imgT + ((size_t)n * p.tiles + jt) * kXTBytes, xb, bar_full);
    }
    if (PASS2 && tid < kGT) {
      const int j = j0 + tid;
      srj[tid] = j < p.k ? make_float2(1.f / __ldg(p.sx + (size_t)n * p.k + j), 1.f / __ldg(p.sz + (size_t)n * p.k + j))
                         : make_float2(0.f, 0.f);
    }
    if (warp == 0) {
      if (elect_one()) {
        if (jt == jt0) mbar_wait(bar_a, 0);
        mbar_wait(bar_full, ph);
        fence_after_sync();
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        for (int s = 0; s < gd.sx; ++s)
          mma_ss(tmem, smem_desc(a0 + s * 4096, 2048, 128), smem_desc(b0 + s * 4096, 2048, 128), idesc, s > 0);
        for (int s = 0; s < gd.sz; ++s)
          mma_ss(tmem + 128, smem_desc(a0 + (gd.sx + s) * 4096, 2048, 128), smem_desc(b0 + (gd.sx + s) * 4096, 2048, 128),
                 idesc, s > 0);
        mma_commit(bar_d);
      }
      __syncwarp();
    }
    mbar_wait(bar_full, ph);                           // x_j rows / norms landed (every thread reads them)
    __syncthreads();                                   // srj visible
    mbar_wait(bar_d, ph);
    fence_after_sync();
    const float2 sJ = snj[kGT];                        // accumulators carry s_I * s_J (powers of two: exact)
    const float cx = -2.f / (sI.x * sJ.x), cz = -2.f / (sI.y * sJ.y);
    const bool edge = (jt == itile) || (j0 + kGT > p.k) || !rows_full;      // CTA-uniform
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      const int col = hf * 64 + cc * 16;
      uint32_t ux[16], uz[16];
      tmem_ld16(lane_base + col, ux);
      tmem_ld16(lane_base + 128 + col, uz);
      tmem_ld_wait();
      float wch[16];
      if (edge) {
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const int jl = col + c, j = j0 + jl;
          wch[c] = 0.f;
          gram_pair<PASS2, true, CHECK>(acc, __uint_as_float(ux[c]), __uint_as_float(uz[c]), ni, snj[jl], cx, cz, ri,
                                        PASS2 ? srj[jl] : make_float2(0.f, 0.f), p.margin, j < p.k, ivalid, j != i, grad,
                                        alpha, wch[c]);
        }
      } else {
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const int jl = col + c;
          wch[c] = 0.f;
          gram_pair<PASS2, false, CHECK>(acc, __uint_as_float(ux[c]), __uint_as_float(uz[c]), ni, snj[jl], cx, cz, ri,
                                         PASS2 ? srj[jl] : make_float2(0.f, 0.f), p.margin, true, true, true, grad,
                                         alpha, wch[c]);
        }
      }
      if (grad) {
        // What (16 consecutive j of this row) -> fp16 hi / lo A-operand columns, written over accumulator columns
        // this warp has already consumed: hi into the x accumulator block, lo into the z block.
        uint32_t whi[8], wlo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          whi[e] = pack_hi(wch[2 * e], wch[2 * e + 1]);
          wlo[e] = pack_lo(wch[2 * e], wch[2 * e + 1], whi[e]);
        }
        const uint32_t wcol = lane_base + hf * 64 + cc * 8;
        tmem_st8(wcol, whi);
        tmem_st8(wcol + 128, wlo);
      }
    }
    if (grad) {
      // G[128 x 32] = What[128 x 128] * C_j[128 x 32] on the tensor pipe (A from TMEM, B = transposed x image)
      tmem_st_wait();
      fence_before_sync();
      __syncthreads();
      if (warp == 0) {
        if (elect_one()) {
          fence_after_sync();
          const uint32_t t0 = smem_u32(sXT);
          const uint32_t idg = idesc_f16(32);
#pragma unroll 1
          for (int sl = 0; sl < 8; ++sl) {             // K slabs of 16 j: packed columns 8*sl (j < 64) or 64 + 8*(sl-4)
            const uint32_t acol = tmem + (sl < 4 ? sl * 8 : 64 + (sl - 4) * 8);
            const uint64_t bh = smem_desc(t0 + sl * 1024, 512, 128), bl = smem_desc(t0 + 8192 + sl * 1024, 512, 128);
            mma_ts(tmem + 32, acol, bh, idg, sl > 0);          // hi * hi
            mma_ts(tmem + 32, acol + 128, bh, idg, 1);         // lo * hi
            mma_ts(tmem + 32, acol, bl, idg, 1);               // hi * lo
          }
          mma_commit(bar_g);
        }
        __syncwarp();
      }
      mbar_wait(bar_g, ph);
      fence_after_sync();
      uint32_t ug[16];
      tmem_ld16(lane_base + 32 + hf * 16, ug);
      tmem_ld_wait();
      const float isj = 1.f / snj[kGT].x;              // the tile's operand scale (power of two)
#pragma unroll
      for (int d = 0; d < 16; ++d) acc.g[d] = fmaf(__uint_as_float(ug[d]), isj, acc.g[d]);
    }
    fence_before_sync();
    __syncthreads();                                   // TMEM and the tile's shared buffers may be overwritten
    ph ^= 1;
  }

  if ((CHECK || grad) && acc.illm < 0.f) atomicOr(p.ill, 1);
  if (!PASS2) {
    sred[hf * kGT + row] = make_float2(acc.rsx, acc.rsz);
    __syncthreads();
    if (tid < kGT) {
      const int ii = itile * kGT + tid;
      if (ii >= p.row0 && ii < iend) {
        const size_t o = ((size_t)js * N + n) * p.k + ii;
        p.psx[o] = sred[tid].x + sred[kGT + tid].x;
        p.psz[o] = sred[tid].y + sred[kGT + tid].y;
      }
    }
  } else {
    float lsum = acc.lsum;
    const float sumW = acc.sumW;
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 16);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 8);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 4);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 2);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 1);
    if (lane == 0) slos[warp] = lsum;
    if (grad) sred[hf * kGT + row] = make_float2(sumW, 0.f);      // What row sums of the two column halves
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < 8; ++w) t += slos[w];
      p.loss_part[((size_t)n * gridDim.y + js) * gridDim.x + blockIdx.x] = t;
    }
    if (grad && ivalid) {
      // grad_i = c_i * sum_j What_ij - sum_j What_ij c_j ; this thread owns dims hf*16 .. hf*16+15 of row i
      const float sw = sred[row].x + sred[kGT + row].x;
      const float ia = 1.f / alpha;
      const size_t off = ((size_t)n * p.k + i) * p.Dx;
      float* dst = p.gpart + (size_t)js * N * p.k * p.Dx + off;
      const float* mu = p.mu + (size_t)n * 64;
#pragma unroll
      for (int d = 0; d < 16; ++d) {
        const int dim = hf * 16 + d;
        if (dim < p.Dx) dst[dim] = fmaf(__ldg(p.x + off + dim) - __ldg(mu + dim), sw, -acc.g[d] * ia);
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc256(tmem);
}

// ---------------------------------------------------------------- host side (called from k2_ndiv.cu)
struct GramPlan { int tiles, tps, JS; };
GramPlan gram_plan(int N, int k) {
  GramPlan g;
  g.tiles = (k + kGT - 1) / kGT;
  long long want = (148 * 2) / ((long long)(N > 0 ? N : 1) * g.tiles);       // one wave of 2 CTAs/SM
  if (want < 1) want = 1;
  if (want > g.tiles) want = g.tiles;
  g.tps = (int)((g.tiles + want - 1) / want);
  g.JS = (g.tiles + g.tps - 1) / g.tps;
  return g;
}

// the fast path covers the shapes of the planning workloads; everything else takes the exact kernels
bool gram_eligible(int N, int k, int Dx, int Dz, int row0) {
  static const bool off = getenv("NDP_NDIV_EXACT") != nullptr;      // diagnostics: force the exact kernels
  if (off) return false;
  return N >= 1 && (long long)N * k >= 1536 && k >= 256 && Dx >= 1 && Dx <= 32 && (Dx % 4) == 0 && Dz >= 1 && Dz <= 32 && (row0 % kGT) == 0;
}

struct GramWs { size_t flag, mu, norms, imgA, imgB, imgT, total; };
static GramWs gram_ws_layout(int N, int k, int Dx, int Dz) {
  const GramDims gd = gram_dims(Dx, Dz);
  const size_t tiles = (size_t)(k + kGT - 1) / kGT;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  GramWs w;
  w.flag = 0;
  w.mu = 256;
  w.norms = w.mu + al((size_t)N * 64 * 4);
  w.imgA = w.norms + al((size_t)N * tiles * kGNS * 8);
  w.imgB = w.imgA + al((size_t)N * tiles * gd.img_bytes);
  w.imgT = w.imgB + al((size_t)N * tiles * gd.img_bytes);
  w.total = w.imgT + al((size_t)N * tiles * kXTBytes);
  return w;
}
size_t gram_ws_bytes(int N, int k, int Dx, int Dz) { return gram_ws_layout(N, k, Dx, Dz).total; }
int gram_loss_parts(int N, int k) { const GramPlan g = gram_plan(N, k); return N * g.JS * g.tiles; }
int gram_js(int N, int k) { return gram_plan(N, k).JS; }

static size_t gram_smem(int Dx, int Dz, bool pass2) {
  return (size_t)2 * gram_dims(Dx, Dz).img_bytes + (pass2 ? (size_t)kXTBytes : 0) + 4096 + 256;
}

// mean -> scale -> operand images + norms; clears the ill-conditioning flag.  ws: gram_ws_bytes()
int gram_prepare(const float* x, const float* z, int N, int k, int Dx, int Dz, void* ws, cudaStream_t s) {
  const GramWs w = gram_ws_layout(N, k, Dx, Dz);
  char* b = reinterpret_cast<char*>(ws);
  const int tiles = (k + kGT - 1) / kGT;
  ndiv_prepack_kernel<<<dim3(tiles, N, 2), 256, 0, s>>>(x, z, k, Dx, Dz, reinterpret_cast<uint8_t*>(b + w.imgA),
                                                     reinterpret_cast<uint8_t*>(b + w.imgB),
                                                     reinterpret_cast<uint8_t*>(b + w.imgT),
                                                     reinterpret_cast<float*>(b + w.mu),
                                                     reinterpret_cast<float2*>(b + w.norms),
                                                     reinterpret_cast<int*>(b + w.flag));
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

int* gram_flag(void* ws) { return reinterpret_cast<int*>(ws); }

static void gram_fill(GramParams& p, int N, int k, int Dx, int Dz, void* ws, int row0, int rows) {
  const GramWs w = gram_ws_layout(N, k, Dx, Dz);
  const GramPlan g = gram_plan(N, k);
  char* b = reinterpret_cast<char*>(ws);
  p.imgA = reinterpret_cast<const uint8_t*>(b + w.imgA);
  p.imgB = reinterpret_cast<const uint8_t*>(b + w.imgB);
  p.norms = reinterpret_cast<const float2*>(b + w.norms);
  p.imgT = reinterpret_cast<const uint8_t*>(b + w.imgT);
  p.mu = reinterpret_cast<const float*>(b + w.mu);
  p.k = k; p.Dx = Dx; p.Dz = Dz; p.tiles = g.tiles; p.tps = g.tps; p.row0 = row0; p.rows = rows;
  p.ill = reinterpret_cast<int*>(b + w.flag);
}

// pass 1 on tensor cores: partial row sums psx/psz [JS][N][k] (JS = gram_js)
int gram_rowsums(int N, int k, int Dx, int Dz, int row0, int rows, void* ws, float* psx, float* psz, cudaStream_t s) {
  GramParams p{};
  gram_fill(p, N, k, Dx, Dz, ws, row0, rows);
  p.psx = psx; p.psz = psz;
  const GramPlan g = gram_plan(N, k);
  const size_t sm = gram_smem(Dx, Dz, false);
  NDP_CUDA(cudaFuncSetAttribute(ndiv_gram_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  ndiv_gram_kernel<false, true><<<dim3((rows + kGT - 1) / kGT, g.JS, N), kGNT, sm, s>>>(p);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// pass 2 on tensor cores: loss partials [gram_loss_parts] and gradient partials gpart [JS][N][k][Dx] (nullable)
int gram_loss_grad(const float* x, const float* sx, const float* sz, int N, int k, int Dx, int Dz, float margin, int row0,
                   int rows, void* ws, float* loss_part, float* gpart, bool check, cudaStream_t s) {
  GramParams p{};
  gram_fill(p, N, k, Dx, Dz, ws, row0, rows);
  p.x = x; p.sx = sx; p.sz = sz; p.margin = margin; p.loss_part = loss_part; p.gpart = gpart;
  const GramPlan g = gram_plan(N, k);
  const size_t sm = gram_smem(Dx, Dz, true);
  const dim3 grid((rows + kGT - 1) / kGT, g.JS, N);
  if (check) {
    NDP_CUDA(cudaFuncSetAttribute(ndiv_gram_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    ndiv_gram_kernel<true, true><<<grid, kGNT, sm, s>>>(p);
  } else {                                             // same pairs were checked by pass 1 of the same call
    NDP_CUDA(cudaFuncSetAttribute(ndiv_gram_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    ndiv_gram_kernel<true, false><<<grid, kGNT, sm, s>>>(p);
  }
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

}  // namespace ndp
