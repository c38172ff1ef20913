// K2 fast path — pairwise distances in Gram form on the 5th-gen tensor cores (north_star:
// ||a||^2 + ||b||^2 - 2ab with the normalisation and hinge epilogue fused), reference diversity.py:8-41.
//
//   d2_ij = n_i + n_j - 2 <c_i, c_j>,   c = v - mean_group(v)   (centering keeps n comparable to d2),
//   <c_i,c_j> by tcgen05.mma kind::f16 on fp16 split operands c*s = hi + lo (22 significand bits):
//   hi.hi + hi.lo + lo.hi, fp32 accumulation in TMEM (one 128x128 tile per matrix and j-tile).
//
// The Gram form loses digits when d2 << n_i + n_j (near-duplicate outputs — the regime the loss exists for).  Every
// pair is therefore tested against d2 >= 2^-6 (n_i + n_j); a pair that fails takes a slow path INSIDE the kernel:
// its distance is recomputed from the fp32 rows as the reference does (sum of squared differences), and in pass 2
// its gradient term What_ij (x_i - x_j) is accumulated directly in fp32 instead of going through the fp16 gradient
// product (where What ~ 1/d would overflow and the two terms of x_i sum(What) - What x would cancel).  Pairs that pass
// have relative error <= 2^-21 * 2^6 / 2 = 1.5e-5 in d (at the threshold; 3e-7 typical), inside north_star's 1e-4.
// The diagonal is exactly 0 by construction.  No fallback launches, no device flag.
//
// Three launches per forward+backward:
//   1. ndiv_prepack_kernel   (4 blocks per 128-row tile): centre, per-tile power-of-two scale, fp16 hi/lo split ->
//      ONE MMA-ready shared-memory image per tile (K-major, no swizzle, 8x16B core matrices, slabs
//      [x_hi | x_lo | z_hi | z_lo]; the same image serves as A and as B operand), centred norms, the transposed x
//      image of the gradient product; clears the reduction tickets.
//   2. ndiv_gram_kernel<false>: row sums.   3. ndiv_gram_kernel<true>: loss + What + What.X.
// One CTA = 128 rows x a range of 128-column tiles, 256 threads (8 epilogue warps: TMEM lane quadrant = warp % 4,
// column half = warp / 4), 2 CTAs / SM (256 TMEM columns each).  The j-tile operands are DOUBLE-BUFFERED in shared
// memory: the 1-D bulk copies (TMA engine) of tile t+2 are issued when tile t retires, so copy latency hides under a
// whole tile of MMA + epilogue.  The column splits' partial results are combined by the LAST CTA of each row tile
// (atomic ticket) in fixed split order — deterministic, and a row-sharded call (blocks aligned to 128) reproduces the
// full call bit for bit; the loss partials are summed the same way by the last CTA of the grid.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "peer.cuh"

namespace ndp {
using namespace tc;

static constexpr int kGT = 128;                 // rows / columns per tile
static constexpr int kGNT = 256;                // threads per CTA
static constexpr float kIllTau = 1.f / 64.f;    // conditioning threshold d2 >= tau * (n_i + n_j)
static constexpr int kXTBytes = 16384;          // per tile: transposed x operand of the gradient product, hi + lo
static constexpr int kGNS = 130;                // 8-byte entries per tile in the norms array (1040 B): floats nx[128] |
                                                // nz[128] | {scale_x, scale_z, 0, 0}  (structure of arrays: two adjacent
                                                // columns' norms load as one 64-bit word for the packed f32x2 epilogue)
static constexpr int kNormBytes = 1056;         // kGNS * 8 rounded up to a multiple of 32 (shared-memory slot)

struct GramDims {
  int gx, gz;          // K=16 slabs per half (hi or lo) of the x / z operands
  int img_bytes;       // bytes of one tile image: [x_hi | x_lo | z_hi | z_lo], 4096 per slab
};
__host__ __device__ inline GramDims gram_dims(int Dx, int Dz) {
  GramDims g;
  g.gx = (Dx + 15) / 16;
  g.gz = (Dz + 15) / 16;
  g.img_bytes = 2 * (g.gx + g.gz) * 4096;
  return g;
}

__device__ __forceinline__ float pick_scale(float m) {   // 2^(9 - e) with m = f * 2^e, f in [0.5, 1); degenerate data -> 1
  if (!(m > 0.f) || !(m < 3e38f)) return 1.f;
  int e;
  frexpf(m, &e);
  int sh = 9 - e;
  sh = sh > 60 ? 60 : (sh < -60 ? -60 : sh);
  return ldexpf(1.f, sh);
}

// byte offset of element (row r, K index kk) inside a tile image whose matrix half starts at slab0
__device__ __forceinline__ uint32_t img_off(int slab0, int r, int kk) {
  return (uint32_t)((slab0 + (kk >> 4)) * 4096 + ((kk >> 3) & 1) * 2048 + (r >> 3) * 128 + (r & 7) * 16 + (kk & 7) * 2);
}

// grid (4 * tiles, N), 256 threads; block = 32 rows of one tile.  img: [N][tiles][img_bytes]; norms: [N][tiles][kGNS]
// float2; imgT: [N][tiles][kXTBytes].  Centre = mean of the group's first min(k,32) rows (any fixed vector is exact —
// distances are translation invariant — it only has to be central enough to keep ||c||^2 comparable to d^2); every
// block recomputes it in the same order, so all tiles agree bit for bit.  Operand scale: a power of two PER TILE
// (max |c| * s in [256, 512)), so the fp16 "lo" halves stay normal numbers whatever the magnitude of the data; the
// main kernel divides the accumulator by s_I * s_J.
__global__ void __launch_bounds__(256)
ndiv_prepack_kernel(const float* __restrict__ x, const float* __restrict__ z, int k, int Dx, int Dz,
                    uint8_t* __restrict__ img, uint8_t* __restrict__ imgT, float* __restrict__ mu_out,
                    float2* __restrict__ norms, unsigned* __restrict__ tickets, int n_tickets) {
  const int n = blockIdx.y, tile = blockIdx.x >> 2, sub = blockIdx.x & 3, tiles = gridDim.x >> 2;
  const int tid = threadIdx.x;
  if (blockIdx.x == 0 && blockIdx.y == 0)
    for (int i = tid; i < n_tickets; i += 256) tickets[i] = 0u;           // consumed by the pass kernels (stream order)
  const GramDims gd = gram_dims(Dx, Dz);
  uint8_t* I = img + ((size_t)n * tiles + tile) * gd.img_bytes;
  __shared__ float mu[64];
  __shared__ float s_part[4][64];
  __shared__ float s_mx[8], s_mz[8];
  __shared__ float s_sc[2];
  {
    const int d = tid & 63, rq = tid >> 6, D = Dx + Dz;
    const int ns = k < 32 ? k : 32;
    float acc = 0.f;
    if (d < D) {
      const float* v = d < Dx ? x + (size_t)n * k * Dx + d : z + (size_t)n * k * Dz + (d - Dx);
      const int st = d < Dx ? Dx : Dz;
#pragma unroll 8
      for (int r = rq; r < ns; r += 4) acc += __ldg(v + (size_t)r * st);
    }
    s_part[rq][d] = acc;
    __syncthreads();
    if (tid < 64) mu[tid] = ((s_part[0][d] + s_part[1][d]) + (s_part[2][d] + s_part[3][d])) / (float)ns;
    __syncthreads();
  }
  const int r0 = tile * kGT, r1 = (r0 + kGT) < k ? (r0 + kGT) : k;
  {
    float mx = 0.f, mz = 0.f;
#pragma unroll 4
    for (int i = tid; i < (r1 - r0) * Dx; i += 256)
      mx = fmaxf(mx, fabsf(__ldg(x + ((size_t)n * k + r0) * Dx + i) - mu[i % Dx]));
#pragma unroll 4
    for (int i = tid; i < (r1 - r0) * Dz; i += 256)
      mz = fmaxf(mz, fabsf(__ldg(z + ((size_t)n * k + r0) * Dz + i) - mu[Dx + i % Dz]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mz = fmaxf(mz, __shfl_xor_sync(0xffffffffu, mz, o));
    }
    if ((tid & 31) == 0) { s_mx[tid >> 5] = mx; s_mz[tid >> 5] = mz; }
    __syncthreads();
    if (tid == 0) {
      float a = 0.f, b = 0.f;
      for (int w = 0; w < 8; ++w) { a = fmaxf(a, s_mx[w]); b = fmaxf(b, s_mz[w]); }
      s_sc[0] = pick_scale(a); s_sc[1] = pick_scale(b);
      if (sub == 0) {
        float* nf = reinterpret_cast<float*>(norms + ((size_t)n * tiles + tile) * kGNS);
        nf[2 * kGT] = s_sc[0]; nf[2 * kGT + 1] = s_sc[1]; nf[2 * kGT + 2] = 0.f; nf[2 * kGT + 3] = 0.f;
      }
    }
    __syncthreads();
  }
  const int gx8 = gd.gx * 2, gz8 = gd.gz * 2;          // 8-dim groups per row (padded)
  for (int it = tid; it < 32 * (gx8 + gz8); it += 256) {
    const int r = sub * 32 + (it & 31), g = it >> 5;
    const bool isx = g < gx8;
    const int g8 = isx ? g : g - gx8;
    const int D = isx ? Dx : Dz;
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
    const int slab_hi = isx ? 0 : 2 * gd.gx, slab_lo = isx ? gd.gx : 2 * gd.gx + gd.gz;
    *reinterpret_cast<uint4*>(I + img_off(slab_hi, r, g8 * 8)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(I + img_off(slab_lo, r, g8 * 8)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
  // B operand of the gradient product  G[i, d] = sum_j What_ij c_j[d]:  rows = d (N = 32), K = j (128), K-major core
  // matrices [j/8][d/8][8 d][8 j]; fp16 hi (8 KB) then lo (8 KB) of the same scaled centred values as above.
  {
    uint8_t* T = imgT + ((size_t)n * tiles + tile) * kXTBytes;
    for (int it = tid; it < 4 * 32; it += 256) {
      const int d = it & 31, jg = sub * 4 + (it >> 5);
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int ra = tile * kGT + jg * 8 + 2 * e;
        const float c0 = (ra < k && d < Dx) ? (__ldg(x + ((size_t)n * k + ra) * Dx + d) - mu[d]) * s_sc[0] : 0.f;
        const float c1 = (ra + 1 < k && d < Dx) ? (__ldg(x + ((size_t)n * k + ra + 1) * Dx + d) - mu[d]) * s_sc[0] : 0.f;
        hi[e] = pack_hi(c0, c1);
        lo[e] = pack_lo(c0, c1, hi[e]);
      }
      const uint32_t off = (uint32_t)((jg * 4 + (d >> 3)) * 128 + (d & 7) * 16);
      *reinterpret_cast<uint4*>(T + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(T + 8192 + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    if (tile == 0 && sub == 0 && tid < 64) mu_out[(size_t)n * 64 + tid] = tid < Dx ? mu[tid] : 0.f;
  }
  if (tid < 32) {                                      // centred squared norms, fixed summation order
    const int row = tile * kGT + sub * 32 + tid;
    float nx = 0.f, nz = 0.f;
    if (row < k) {
      const float* xr = x + ((size_t)n * k + row) * Dx;
      const float* zr = z + ((size_t)n * k + row) * Dz;
      for (int d = 0; d < Dx; ++d) { const float c = __ldg(xr + d) - mu[d]; nx = fmaf(c, c, nx); }
      for (int d = 0; d < Dz; ++d) { const float c = __ldg(zr + d) - mu[Dx + d]; nz = fmaf(c, c, nz); }
    }
    float* nf = reinterpret_cast<float*>(norms + ((size_t)n * tiles + tile) * kGNS);
    nf[sub * 32 + tid] = nx;
    nf[kGT + sub * 32 + tid] = nz;
  }
}

// MUFU.SQRT (<= 1 ulp-class error, same order as the Gram form's own rounding); IEEE sqrtf runs at 9.7 /clk/SM against
// 15.8 for the approximation (tools/fp32_probe.cu) and the epilogue is MUFU-bound.  The exact kernels keep sqrtf.
__device__ __forceinline__ float sqrt_approx(float v) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }

struct GramParams {
  const uint8_t* img;
  const float2* norms;
  const uint8_t* imgT;         // transposed x operand images (gradient product, pass 2)
  const float* mu;             // [N][64] centre of the x features
  const float* x;              // original rows (slow path; pass 2 gradient)
  const float* z;
  const float* sx;             // row sums (pass 2 input)
  const float* sz;
  int k, Dx, Dz, tiles, tps, JS;
  int row0, rows;              // rows owned by this call (row0 is a multiple of 128)
  float margin;
  float* psx;                  // pass 1: partial row sums [JS][N][k] (JS > 1)
  float* psz;
  float* out_sx;               // pass 1: final row sums [N][k]
  float* out_sz;
  float* loss_part;            // pass 2: [N][JS][row tiles of the call]
  float* loss;                 // pass 2: final loss
  float* gpart;                // pass 2: [JS][N][k][Dx] (JS > 1)
  float* grad;                 // pass 2: final gradient [N][k][Dx] (nullable: loss only)
  unsigned* tickets;           // [2 * N * tiles + 2]: pass-1 row-tile tickets, pass-2 row-tile tickets, grid tickets of
                               // pass 2 and pass 1
  // row-sharded call over a peer window (peer.cuh): the LAST CTA of the grid publishes `sig_epoch` in flag slot
  // `sig_slot` of every rank once this rank's results are written (pw.world == 0: single-GPU call, no signal)
  PeerWin pw;
  int sig_slot;
  unsigned sig_epoch;
};

// exact squared distance of two fp32 rows (the reference's own form: sum of squared differences)
__device__ __noinline__ float exact_d2(const float* __restrict__ a, const float* __restrict__ b, int D) {
  float s = 0.f;
  for (int d = 0; d < D; ++d) { const float df = __ldg(a + d) - __ldg(b + d); s = fmaf(df, df, s); }
  return s;
}

// packed fp32 pairs (one issue slot for two lanes' worth of work: the epilogue is issue-bound)
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void up2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float rsqrt_approx(float v) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }

static constexpr float kTinyD2 = 1e-30f;              // pass 2 clamps d^2 here so that rsqrt stays finite

// per-thread state of the epilogue (one row i of the tile, one half of its columns)
template <bool PASS2>
struct RowState {
  float rsx, rsz;                                     // pass 1: row sums
  float lsum, sumW, wmax;                             // pass 2: loss, sum_j What_ij, max |What * alpha| of the chunk
  float* gdir;                                        // pass 2: directly accumulated gradient terms (all 32 dims; an
                                                      // array in LOCAL memory: only the rare slow path touches it)
  bool any_direct;
};

// 16 consecutive columns [col, col+16) of row i: squared distances from the two Gram accumulators (+ the exact slow
// path for ill-conditioned pairs), then row sums (pass 1) or hinge / loss / What (pass 2).  EDGE: the tile contains
// the diagonal, padded columns or rows outside the caller's block (CTA-uniform).  wch: What of the 16 pairs.
template <bool PASS2, bool EDGE>
__device__ __forceinline__ void gram_chunk(RowState<PASS2>& st, const uint32_t (&ux)[16], const uint32_t (&uz)[16],
                                           const float* __restrict__ snx, const float* __restrict__ snz,
                                           const float* __restrict__ srx, const float* __restrict__ srz,
                                           int col, float nix, float niz, float cx, float cz, float rix, float mriz,
                                           float alpha, bool grad, bool ivalid, int i, int j0, int k,
                                           const float* __restrict__ xi, const float* __restrict__ zi,
                                           const float* __restrict__ xg, const float* __restrict__ zg, int Dx, int Dz,
                                           float (&wch)[16]) {
  float d2x[16], d2z[16];
  float illm = 1.f;                                    // running min of d2 - tau (n_i + n_j) over the chunk
  {
    const u64 nix2 = pk2(nix, nix), niz2 = pk2(niz, niz), cx2 = pk2(cx, cx), cz2 = pk2(cz, cz);
    const u64 ntau2 = pk2(-kIllTau, -kIllTau);
#pragma unroll
    for (int pq = 0; pq < 8; ++pq) {
      const u64 sxn = add2(nix2, *reinterpret_cast<const u64*>(snx + col + 2 * pq));
      const u64 szn = add2(niz2, *reinterpret_cast<const u64*>(snz + col + 2 * pq));
      const u64 vx = fma2(cx2, pk2(__uint_as_float(ux[2 * pq]), __uint_as_float(ux[2 * pq + 1])), sxn);
      const u64 vz = fma2(cz2, pk2(__uint_as_float(uz[2 * pq]), __uint_as_float(uz[2 * pq + 1])), szn);
      float a0, a1, b0, b1, e0, e1, f0, f1;
      up2(vx, a0, a1); up2(vz, b0, b1);
      up2(fma2(ntau2, sxn, vx), e0, e1); up2(fma2(ntau2, szn, vz), f0, f1);
      if (EDGE) {                                      // only live pairs take part in the conditioning test
        const int j = j0 + col + 2 * pq;
        const bool l0 = ivalid && j < k && j != i, l1 = ivalid && j + 1 < k && j + 1 != i;
        e0 = l0 ? e0 : 1.f; f0 = l0 ? f0 : 1.f; e1 = l1 ? e1 : 1.f; f1 = l1 ? f1 : 1.f;
        if (j == i) { a0 = 0.f; b0 = 0.f; }            // the diagonal is exactly 0
        if (j + 1 == i) { a1 = 0.f; b1 = 0.f; }
      }
      illm = fminf(illm, fminf(fminf(e0, e1), fminf(f0, f1)));
      if (EDGE) {                                      // diagonal zeros / padded columns must stay sqrt- and rsqrt-safe
        const float lo = PASS2 ? kTinyD2 : 0.f;
        d2x[2 * pq] = fmaxf(a0, lo); d2x[2 * pq + 1] = fmaxf(a1, lo);
        d2z[2 * pq] = fmaxf(b0, lo); d2z[2 * pq + 1] = fmaxf(b1, lo);
      } else {
        // interior tile: no clamp.  A value that is not comfortably positive (d2 <= tau (n_i + n_j), negative
        // round-off and NaN included) fails the conditioning test above and is REPLACED by the exact, clamped
        // distance in the slow path below before any sqrt / rsqrt sees it
        d2x[2 * pq] = a0; d2x[2 * pq + 1] = a1;
        d2z[2 * pq] = b0; d2z[2 * pq + 1] = b1;
      }
    }
  }
  uint32_t direct = 0;                                 // pairs whose gradient term is accumulated directly
  if (!(illm > 0.f)) {                                 // rare: near-duplicate rows -> the reference's own difference form
#pragma unroll 1
    for (int c = 0; c < 16; ++c) {
      const int j = j0 + col + c;
      if (EDGE && !(ivalid && j < k && j != i)) continue;
      float vx = 0.f, vz = 0.f;
#pragma unroll
      for (int c2 = 0; c2 < 16; ++c2)
        if (c2 == c) { vx = d2x[c2]; vz = d2z[c2]; }
      const float sxn = nix + snx[col + c], szn = niz + snz[col + c];
      if (vx > kIllTau * sxn && vz > kIllTau * szn) continue;
      const float ex = exact_d2(xi, xg + (size_t)j * Dx, Dx), ez = exact_d2(zi, zg + (size_t)j * Dz, Dz);
#pragma unroll
      for (int c2 = 0; c2 < 16; ++c2)
        if (c2 == c) { d2x[c2] = PASS2 ? fmaxf(ex, kTinyD2) : ex; d2z[c2] = PASS2 ? fmaxf(ez, kTinyD2) : ez; }
      direct |= 1u << c;
    }
  }
  if (!PASS2) {
    u64 rs2x = pk2(0.f, 0.f), rs2z = pk2(0.f, 0.f);
#pragma unroll
    for (int pq = 0; pq < 8; ++pq) {
      float ax = sqrt_approx(d2x[2 * pq]), bx = sqrt_approx(d2x[2 * pq + 1]);
      float az = sqrt_approx(d2z[2 * pq]), bz = sqrt_approx(d2z[2 * pq + 1]);
      if (EDGE) {
        const int j = j0 + col + 2 * pq;
        if (j >= k) { ax = 0.f; az = 0.f; }
        if (j + 1 >= k) { bx = 0.f; bz = 0.f; }
      }
      rs2x = add2(rs2x, pk2(ax, bx));
      rs2z = add2(rs2z, pk2(az, bz));
    }
    float a, b;
    up2(rs2x, a, b); st.rsx += a + b;
    up2(rs2z, a, b); st.rsz += a + b;
  } else {
    const u64 rix2 = pk2(rix, rix), mriz2 = pk2(mriz, mriz), nalpha2 = pk2(-alpha, -alpha);
    float wm = 0.f;
    u64 ls2 = pk2(0.f, 0.f), sw2 = pk2(0.f, 0.f);
#pragma unroll
    for (int pq = 0; pq < 8; ++pq) {
      // d = d2 * rsqrt(d2); 1/d = rsqrt(d2): one MUFU per matrix, no reciprocal
      const float r0 = rsqrt_approx(d2x[2 * pq]), r1 = rsqrt_approx(d2x[2 * pq + 1]);
      const u64 ddx = mul2(pk2(d2x[2 * pq], d2x[2 * pq + 1]), pk2(r0, r1));
      const u64 ddz = mul2(pk2(d2z[2 * pq], d2z[2 * pq + 1]),
                           pk2(rsqrt_approx(d2z[2 * pq]), rsqrt_approx(d2z[2 * pq + 1])));
      // hinge arguments h_ij (row i's normalisers) and h_ji (row j's): margin dz / sz - dx / sx
      const u64 rjx = *reinterpret_cast<const u64*>(srx + col + 2 * pq);
      const u64 hij = fma2(ddz, mriz2, mul2(ddx, pk2(-rix, -rix)));
      const u64 hji = fma2(ddz, *reinterpret_cast<const u64*>(srz + col + 2 * pq), mul2(ddx, rjx ^ 0x8000000080000000ull));
      float h0, h1, g0, g1, q0, q1;
      up2(hij, h0, h1); up2(hji, g0, g1); up2(rjx, q0, q1);
      bool ok0 = true, ok1 = true;
      if (EDGE) {
        const int j = j0 + col + 2 * pq;
        ok0 = ivalid && j < k && j != i; ok1 = ivalid && j + 1 < k && j + 1 != i;
      }
      const float l0 = (ok0 && h0 > 0.f) ? h0 : 0.f, l1 = (ok1 && h1 > 0.f) ? h1 : 0.f;
      ls2 = add2(ls2, pk2(l0, l1));
      if (grad) {
        float w0 = (h0 > 0.f ? rix : 0.f) + (g0 > 0.f ? q0 : 0.f);
        float w1 = (h1 > 0.f ? rix : 0.f) + (g1 > 0.f ? q1 : 0.f);
        w0 = ok0 ? -w0 * r0 : 0.f;                     // What = W / d
        w1 = ok1 ? -w1 * r1 : 0.f;
        wch[2 * pq] = w0; wch[2 * pq + 1] = w1;
        wm = fmaxf(wm, fmaxf(fabsf(w0), fabsf(w1)));
      }
    }
    float a, b;
    up2(ls2, a, b); st.lsum += a + b;
    if (grad) {
      if (wm * alpha >= 60000.f) {                     // would not fit the fp16 operand of the gradient product
#pragma unroll
        for (int c = 0; c < 16; ++c)
          if (fabsf(wch[c]) * alpha >= 60000.f) direct |= 1u << c;
      }
      if (direct) {                                    // rare: What_ij (x_i - x_j) in fp32, kept out of the fp16 product
        st.any_direct = true;
#pragma unroll 1
        for (uint32_t m = direct; m; m &= m - 1) {
          const int c = __ffs(m) - 1, j = j0 + col + c;
          float w = 0.f;
#pragma unroll
          for (int c2 = 0; c2 < 16; ++c2)
            if (c2 == c) { w = wch[c2]; wch[c2] = 0.f; }
          if (!(fabsf(w) < 1e29f)) w = 0.f;            // d clamped at 1e-15 (exact duplicates): subgradient 0, as torch.norm
          const float* xj = xg + (size_t)j * Dx;
#pragma unroll 1
          for (int d = 0; d < Dx; ++d) st.gdir[d] = fmaf(w, __ldg(xi + d) - __ldg(xj + d), st.gdir[d]);
        }
      }
#pragma unroll
      for (int pq = 0; pq < 8; ++pq) sw2 = add2(sw2, pk2(wch[2 * pq], wch[2 * pq + 1]));
      up2(sw2, a, b); st.sumW += a + b;
    }
  }
}

#ifdef NDP_TC_PROF
// dev build: per-CTA phase clocks of thread 0, summed over CTAs ([pass][0] CTAs, [1] prologue, [2] first operands
// landed, [3] main loop, [4] per-CTA results + row-tile reduction, [5] grid tail) and the max over CTAs of the total [6]
__device__ unsigned long long g_gram_prof[2][8];
#define GPROF(k_) do { if (tid == 0) { const long long c_ = clock64(); atomicAdd(&g_gram_prof[PASS2][k_], (unsigned long long)(c_ - gp_last)); gp_last = c_; } } while (0)
#else
#define GPROF(k_) do { } while (0)
#endif

// grid (row tiles of the call, JS, N), 256 threads, 2 CTAs / SM
template <bool PASS2>
__global__ void __launch_bounds__(kGNT, 2) ndiv_gram_kernel(const GramParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const GramDims gd = gram_dims(p.Dx, p.Dz);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = blockIdx.z, js = blockIdx.y, N = gridDim.z;
  const int itile = p.row0 / kGT + blockIdx.x;
  const int iend = (p.row0 + p.rows) < p.k ? (p.row0 + p.rows) : p.k;
  const int jt0 = js * p.tps, jt1 = (jt0 + p.tps) < p.tiles ? (jt0 + p.tps) : p.tiles;
  const int ntl = jt1 - jt0;
  const bool grad = PASS2 && p.grad != nullptr;
#ifdef NDP_TC_PROF
  const long long gp_begin = clock64();
  long long gp_last = gp_begin;
  if (tid == 0) atomicAdd(&g_gram_prof[PASS2][0], 1ull);
#endif

  // shared memory: A image | 2 x { B image, norms, [XT] } | srx, srz | sred | slos | barriers
  const uint32_t buf_bytes = (uint32_t)gd.img_bytes + kNormBytes + (PASS2 ? (uint32_t)kXTBytes : 0u);
  uint8_t* sA = smem;
  uint8_t* sBuf = smem + gd.img_bytes;
  float* srx = reinterpret_cast<float*>(sBuf + 2 * buf_bytes);               // 1/sx_j            [128]
  float* srz = srx + kGT;                                                    // margin / sz_j     [128]
  float2* sred = reinterpret_cast<float2*>(srz + kGT);                       // [2 halves][128 rows]
  float* slos = reinterpret_cast<float*>(sred + 2 * kGT);                    // [8 warps]
  uint8_t* sbar = reinterpret_cast<uint8_t*>(slos + 8);
  const uint32_t bar_a = smem_u32(sbar), bar_full0 = bar_a + 8, bar_d = bar_a + 24, bar_g = bar_a + 32;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sbar + 48);
  volatile unsigned* s_ticket = reinterpret_cast<volatile unsigned*>(sbar + 56);

  if (tid == 0) {
    mbar_init(bar_a, 1); mbar_init(bar_full0, 1); mbar_init(bar_full0 + 8, 1); mbar_init(bar_d, 1); mbar_init(bar_g, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc256(smem_u32((const void*)tmem_slot));
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  auto issue_tile = [&](int t) {                       // bulk copies of j-tile jt0 + t into buffer t & 1 (one thread)
    const int jt = jt0 + t;
    uint8_t* b = sBuf + (t & 1) * buf_bytes;
    const uint32_t bar = bar_full0 + 8 * (t & 1);
    const uint32_t xb = grad ? (uint32_t)kXTBytes : 0u;
    mbar_expect_tx(bar, (uint32_t)gd.img_bytes + kGNS * 8 + xb);
    bulk_g2s(smem_u32(b), p.img + ((size_t)n * p.tiles + jt) * gd.img_bytes, (uint32_t)gd.img_bytes, bar);
    bulk_g2s(smem_u32(b + gd.img_bytes), p.norms + ((size_t)n * p.tiles + jt) * kGNS, kGNS * 8, bar);
    if (grad) bulk_g2s(smem_u32(b + gd.img_bytes + kNormBytes), p.imgT + ((size_t)n * p.tiles + jt) * kXTBytes, xb, bar);
  };
  if (tid == 0) {
    mbar_expect_tx(bar_a, (uint32_t)gd.img_bytes);
    bulk_g2s(smem_u32(sA), p.img + ((size_t)n * p.tiles + itile) * gd.img_bytes, (uint32_t)gd.img_bytes, bar_a);
    issue_tile(0);
    if (ntl > 1) issue_tile(1);
  }

  // epilogue role: TMEM lane quadrant q, column half hf; this thread owns row `row` of the tile
  const int q = warp & 3, hf = warp >> 2, row = q * 32 + lane;
  const int i = itile * kGT + row;
  const bool ivalid = i >= p.row0 && i < iend;
  const float* nfi = reinterpret_cast<const float*>(p.norms + ((size_t)n * p.tiles + itile) * kGNS);
  const float nix = __ldg(nfi + row), niz = __ldg(nfi + kGT + row);
  const float sIx = __ldg(nfi + 2 * kGT), sIz = __ldg(nfi + 2 * kGT + 1);      // operand scales of the i tile
  float rix = 0.f, mriz = 0.f, nanprobe = 0.f;
  float alpha = 1.f;                                   // power of two ~ sx_i^2 / k: What * alpha = O(mean d / d)
  if (PASS2 && i < p.k) {
    const float sxi = __ldg(p.sx + (size_t)n * p.k + i), szi = __ldg(p.sz + (size_t)n * p.k + i);
    rix = 1.f / sxi; mriz = p.margin / szi;
    nanprobe = 0.f * (sxi + szi);                      // NaN / inf normalisers make the loss NaN like the reference
    const float t = sxi * sxi / (float)p.k;
    if (t > 0.f && t < 3e38f) { int e; frexpf(t, &e); e = e > 100 ? 100 : (e < -100 ? -100 : e); alpha = ldexpf(1.f, e - 1); }
  }
  const float* xg = p.x + (size_t)n * p.k * p.Dx;
  const float* zg = p.z + (size_t)n * p.k * p.Dz;
  const float* xi = xg + (size_t)(i < p.k ? i : 0) * p.Dx;
  const float* zi = zg + (size_t)(i < p.k ? i : 0) * p.Dz;
  const uint32_t idesc = idesc_f16(128);
  const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);

  RowState<PASS2> st;
  st.rsx = st.rsz = st.lsum = st.sumW = st.wmax = 0.f;
  st.any_direct = false;
  float gdir_local[32];
  st.gdir = gdir_local;
  if (PASS2) {
#pragma unroll 1
    for (int d = 0; d < 32; ++d) gdir_local[d] = 0.f;
  }
  float g[16];                                         // gradient accumulators: dims hf*16 .. hf*16+15 of this row
#pragma unroll
  for (int d = 0; d < 16; ++d) g[d] = 0.f;
  if (PASS2 && hf == 0 && ivalid) st.lsum = nanprobe;
  const bool rows_full = itile * kGT >= p.row0 && itile * kGT + kGT <= iend;
  GPROF(1);

  for (int t = 0; t < ntl; ++t) {
    const int jt = jt0 + t, j0 = jt * kGT;
    const uint32_t ph = (uint32_t)(t & 1), bar_full = bar_full0 + 8 * (t & 1), phf = (uint32_t)((t >> 1) & 1);
    uint8_t* sB = sBuf + (t & 1) * buf_bytes;
    const float* snx = reinterpret_cast<const float*>(sB + gd.img_bytes);
    const float* snz = snx + kGT;
    if (PASS2 && tid < kGT) {
      const int j = j0 + tid;
      srx[tid] = j < p.k ? 1.f / __ldg(p.sx + (size_t)n * p.k + j) : 0.f;
      srz[tid] = j < p.k ? p.margin / __ldg(p.sz + (size_t)n * p.k + j) : 0.f;
    }
    if (warp == 0) {
      if (elect_one()) {
        if (t == 0) mbar_wait(bar_a, 0);
        mbar_wait(bar_full, phf);
        fence_after_sync();
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        // x: slabs [0,gx) hi, [gx,2gx) lo;  z: [2gx, 2gx+gz) hi, [2gx+gz, 2gx+2gz) lo
        for (int s = 0; s < gd.gx; ++s) {
          const uint64_t ah = smem_desc(a0 + s * 4096, 2048, 128), al = smem_desc(a0 + (gd.gx + s) * 4096, 2048, 128);
          const uint64_t bh = smem_desc(b0 + s * 4096, 2048, 128), bl = smem_desc(b0 + (gd.gx + s) * 4096, 2048, 128);
          mma_ss(tmem, ah, bh, idesc, s > 0);
          mma_ss(tmem, ah, bl, idesc, 1);
          mma_ss(tmem, al, bh, idesc, 1);
        }
        const uint32_t az = a0 + 2 * gd.gx * 4096, bz = b0 + 2 * gd.gx * 4096;
        for (int s = 0; s < gd.gz; ++s) {
          const uint64_t ah = smem_desc(az + s * 4096, 2048, 128), al = smem_desc(az + (gd.gz + s) * 4096, 2048, 128);
          const uint64_t bh = smem_desc(bz + s * 4096, 2048, 128), bl = smem_desc(bz + (gd.gz + s) * 4096, 2048, 128);
          mma_ss(tmem + 128, ah, bh, idesc, s > 0);
          mma_ss(tmem + 128, ah, bl, idesc, 1);
          mma_ss(tmem + 128, al, bh, idesc, 1);
        }
        mma_commit(bar_d);
      }
      __syncwarp();
    }
    mbar_wait(bar_full, phf);                          // norms landed (every thread reads them)
    if (t == 0) GPROF(2);
    __syncthreads();                                   // srx / srz visible
    mbar_wait(bar_d, ph);
    fence_after_sync();
    const float sJx = snx[2 * kGT], sJz = snx[2 * kGT + 1];   // accumulators carry s_I * s_J (powers of two: exact)
    const float cx = -2.f / (sIx * sJx), cz = -2.f / (sIz * sJz);
    const bool edge = (jt == itile) || (j0 + kGT > p.k) || !rows_full;      // CTA-uniform
#pragma unroll 1
    for (int cc = 0; cc < 4; ++cc) {
      const int col = hf * 64 + cc * 16;
      uint32_t ux[16], uz[16];
      tmem_ld16(lane_base + col, ux);
      tmem_ld16(lane_base + 128 + col, uz);
      tmem_ld_wait();
      float wch[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) wch[c] = 0.f;
      if (edge)
        gram_chunk<PASS2, true>(st, ux, uz, snx, snz, srx, srz, col, nix, niz, cx, cz, rix, mriz, alpha, grad, ivalid, i,
                                j0, p.k, xi, zi, xg, zg, p.Dx, p.Dz, wch);
      else
        gram_chunk<PASS2, false>(st, ux, uz, snx, snz, srx, srz, col, nix, niz, cx, cz, rix, mriz, alpha, grad, true, i,
                                 j0, p.k, xi, zi, xg, zg, p.Dx, p.Dz, wch);
      if (grad) {
        // What (16 consecutive j of this row) -> fp16 hi / lo A-operand columns, written over accumulator columns
        // this warp has already consumed: hi into the x accumulator block, lo into the z block.
        uint32_t whi[8], wlo[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float w0 = wch[2 * e] * alpha, w1 = wch[2 * e + 1] * alpha;
          whi[e] = pack_hi(w0, w1);
          wlo[e] = pack_lo(w0, w1, whi[e]);
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
          const uint32_t t0 = smem_u32(sB + gd.img_bytes + kNormBytes);
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
      const float isj = 1.f / sJx;                     // the tile's operand scale (power of two)
#pragma unroll
      for (int d = 0; d < 16; ++d) g[d] = fmaf(__uint_as_float(ug[d]), isj, g[d]);
    }
    fence_before_sync();
    __syncthreads();                                   // TMEM and this tile's shared buffer may be overwritten
    if (tid == 0 && t + 2 < ntl) issue_tile(t + 2);    // lands under the whole of tile t+1
  }

  GPROF(3);
  // ---- per-CTA results, then the LAST CTA of the row tile combines the column splits in fixed order
  const size_t tk_row = (size_t)(PASS2 ? N * p.tiles : 0) + (size_t)n * p.tiles + itile;
  if (!PASS2) {
    sred[hf * kGT + row] = make_float2(st.rsx, st.rsz);
    __syncthreads();
    if (tid < kGT) {
      const int ii = itile * kGT + tid;
      if (ii >= p.row0 && ii < iend) {
        const float a = sred[tid].x + sred[kGT + tid].x, b = sred[tid].y + sred[kGT + tid].y;
        if (p.JS == 1) { p.out_sx[(size_t)n * p.k + ii] = a; p.out_sz[(size_t)n * p.k + ii] = b; }
        else { const size_t o = ((size_t)js * N + n) * p.k + ii; p.psx[o] = a; p.psz[o] = b; }
      }
    }
    if (p.JS > 1) {
      __threadfence();
      __syncthreads();
      if (tid == 0) *s_ticket = atomicAdd(p.tickets + tk_row, 1u);
      __syncthreads();
      if (*s_ticket == (unsigned)(p.JS - 1)) {
        __threadfence();
        if (tid < kGT) {
          const int ii = itile * kGT + tid;
          if (ii >= p.row0 && ii < iend) {
            float a = 0.f, b = 0.f;
#pragma unroll 8
            for (int s = 0; s < p.JS; ++s) {
              const size_t o = ((size_t)s * N + n) * p.k + ii;
              a += __ldcg(p.psx + o); b += __ldcg(p.psz + o);
            }
            p.out_sx[(size_t)n * p.k + ii] = a; p.out_sz[(size_t)n * p.k + ii] = b;
          }
        }
      }
    }
    GPROF(4);
    if (p.pw.world > 0) {                                  // sharded: the grid's last CTA signals every rank
      __threadfence();
      __syncthreads();
      if (tid == 0) s_ticket[1] = atomicAdd(p.tickets + 2 * (size_t)N * p.tiles + 1, 1u);
      __syncthreads();
      if (s_ticket[1] == gridDim.x * gridDim.y * gridDim.z - 1 && warp == 0) peer_signal_all(p.pw, p.sig_slot, p.sig_epoch);
    }
  } else {
    float lsum = st.lsum;
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 16);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 8);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 4);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 2);
    lsum = lsum + __shfl_xor_sync(0xffffffffu, lsum, 1);
    if (lane == 0) slos[warp] = lsum;
    if (grad) sred[hf * kGT + row] = make_float2(st.sumW, 0.f);   // What row sums of the two column halves
    // the directly accumulated terms cover all 32 dims of this thread's COLUMNS; the two column halves of a row
    // meet through the (now dead) operand buffers: gx[half][row][32]
    float* gxch = reinterpret_cast<float*>(sBuf);
    const int any_direct = grad ? __syncthreads_or(st.any_direct ? 1 : 0) : (__syncthreads(), 0);
    if (any_direct) {
#pragma unroll 1
      for (int d = 0; d < 32; ++d) gxch[(hf * kGT + row) * 33 + d] = st.gdir[d];
      __syncthreads();
    }
    const int n_parts = N * (int)gridDim.y * (int)gridDim.x;
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < 8; ++w) t += slos[w];
      p.loss_part[((size_t)n * gridDim.y + js) * gridDim.x + blockIdx.x] = t;
    }
    if (grad) {
      // grad_i = c_i * sum_j What_ij - sum_j What_ij c_j (+ the directly accumulated terms).  Each thread holds dims
      // hf*16 .. hf*16+15 of ITS row: written to global memory from here, a warp instruction touched 32 different
      // lines (row stride 128 B) for the x loads and the stores alike.  The tile goes through shared memory instead
      // (behind the exchange area of the direct terms) and leaves with coalesced 128-bit loads / stores.
      float* gout = gxch + 2 * kGT * 33;                        // [128 rows][33]
      const float ia = 1.f / alpha;
#pragma unroll
      for (int d = 0; d < 16; ++d) {
        const int dim = hf * 16 + d;
        float v = -g[d] * ia;
        if (any_direct) v += gxch[row * 33 + dim] + gxch[(kGT + row) * 33 + dim];
        gout[row * 33 + dim] = v;
      }
      __syncthreads();
      const int r0c = itile * kGT;
      float* dstt = (p.JS == 1 ? p.grad : p.gpart + (size_t)js * N * p.k * p.Dx) + ((size_t)n * p.k + r0c) * p.Dx;
      const float* xt = p.x + ((size_t)n * p.k + r0c) * p.Dx;
      const float* mu = p.mu + (size_t)n * 64;
      const bool vec = (reinterpret_cast<uintptr_t>(dstt) & 15) == 0 && (reinterpret_cast<uintptr_t>(xt) & 15) == 0;
      for (int e4 = tid; e4 < kGT * p.Dx / 4; e4 += kGNT) {     // Dx % 4 == 0
        const int rr = (e4 * 4) / p.Dx, dim = (e4 * 4) % p.Dx, ii = r0c + rr;
        if (ii < p.row0 || ii >= iend) continue;
        const float sw = sred[rr].x + sred[kGT + rr].x;
        const float* go = gout + rr * 33 + dim;
        float4 xv;
        if (vec) xv = __ldg(reinterpret_cast<const float4*>(xt) + e4);
        else { xv.x = __ldg(xt + e4 * 4); xv.y = __ldg(xt + e4 * 4 + 1); xv.z = __ldg(xt + e4 * 4 + 2); xv.w = __ldg(xt + e4 * 4 + 3); }
        float4 o;
        o.x = fmaf(xv.x - __ldg(mu + dim), sw, go[0]);
        o.y = fmaf(xv.y - __ldg(mu + dim + 1), sw, go[1]);
        o.z = fmaf(xv.z - __ldg(mu + dim + 2), sw, go[2]);
        o.w = fmaf(xv.w - __ldg(mu + dim + 3), sw, go[3]);
        if (vec) reinterpret_cast<float4*>(dstt)[e4] = o;
        else { dstt[e4 * 4] = o.x; dstt[e4 * 4 + 1] = o.y; dstt[e4 * 4 + 2] = o.z; dstt[e4 * 4 + 3] = o.w; }
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket[0] = (grad && p.JS > 1) ? atomicAdd(p.tickets + tk_row, 1u) : 0u;
    __syncthreads();
    if (grad && p.JS > 1 && s_ticket[0] == (unsigned)(p.JS - 1)) {
      __threadfence();
      const int r0c = itile * kGT;
      const size_t split = (size_t)N * p.k * p.Dx;                    // floats between two column splits' partials
      const size_t tile_off = ((size_t)n * p.k + r0c) * p.Dx;
      const float* __restrict__ gp = p.gpart + tile_off;
      float* __restrict__ go = p.grad + tile_off;
      if ((reinterpret_cast<uintptr_t>(go) & 15) == 0) {
        // 128-bit loads, eight splits in flight per element group: this tail is latency-bound (JS x 16 KB read by ONE
        // CTA while the others have left) and was the largest fixed cost of the pass (profiles/r02_summary.md)
        const int nv = kGT * p.Dx / 4;                                // Dx % 4 == 0
#pragma unroll 2
        for (int e4 = tid; e4 < nv; e4 += kGNT) {
          const int ii = r0c + (e4 * 4) / p.Dx;
          if (ii < p.row0 || ii >= iend) continue;
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int s0 = 0; s0 < p.JS; s0 += 8) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
              if (s0 + u < p.JS) v[u] = __ldcg(reinterpret_cast<const float4*>(gp + (size_t)(s0 + u) * split) + e4);
#pragma unroll
            for (int u = 0; u < 8; ++u)                               // fixed split order
              if (s0 + u < p.JS) { a.x += v[u].x; a.y += v[u].y; a.z += v[u].z; a.w += v[u].w; }
          }
          reinterpret_cast<float4*>(go)[e4] = a;
        }
      } else {
        for (int e = tid; e < kGT * p.Dx; e += kGNT) {
          const int ii = r0c + e / p.Dx;
          if (ii >= p.row0 && ii < iend) {
            float a = 0.f;
#pragma unroll 8
            for (int s = 0; s < p.JS; ++s) a += __ldcg(gp + (size_t)s * split + e);   // fixed split order
            go[e] = a;
          }
        }
      }
      __threadfence();
    }
    GPROF(4);
    // grid ticket AFTER this CTA's last global write (the combined gradient rows above), so that whoever draws the
    // last ticket may publish the whole call's results
    __syncthreads();
    if (tid == 0) s_ticket[1] = atomicAdd(p.tickets + 2 * (size_t)N * p.tiles, 1u);
    __syncthreads();
    if (s_ticket[1] == (unsigned)(n_parts - 1)) {        // last CTA of the grid: the loss, fixed summation order
      __threadfence();
      float t = 0.f;
      for (int e = tid; e < n_parts; e += kGNT) t += __ldcg(p.loss_part + e);
      float* sf = reinterpret_cast<float*>(sred);
      __syncthreads();
      sf[tid] = t;
      __syncthreads();
      for (int o = kGNT / 2; o > 0; o >>= 1) {
        if (tid < o) sf[tid] += sf[tid + o];
        __syncthreads();
      }
      if (tid == 0) p.loss[0] = sf[0];
      if (p.pw.world > 0) {
        __syncthreads();
        if (warp == 0) peer_signal_all(p.pw, p.sig_slot, p.sig_epoch);
      }
    }
  }
  GPROF(5);
#ifdef NDP_TC_PROF
  if (tid == 0) atomicMax(&g_gram_prof[PASS2][6], (unsigned long long)(clock64() - gp_begin));
#endif
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

struct GramWs { size_t tickets, mu, norms, img, imgT, total; };
static GramWs gram_ws_layout(int N, int k, int Dx, int Dz) {
  const GramDims gd = gram_dims(Dx, Dz);
  const size_t tiles = (size_t)(k + kGT - 1) / kGT;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  GramWs w;
  w.tickets = 0;
  w.mu = al((2 * (size_t)N * tiles + 2) * sizeof(unsigned));
  w.norms = w.mu + al((size_t)N * 64 * 4);
  w.img = w.norms + al((size_t)N * tiles * kGNS * 8);
  w.imgT = w.img + al((size_t)N * tiles * gd.img_bytes);
  w.total = w.imgT + al((size_t)N * tiles * kXTBytes);
  return w;
}
size_t gram_ws_bytes(int N, int k, int Dx, int Dz) { return gram_ws_layout(N, k, Dx, Dz).total; }
int gram_loss_parts(int N, int k) { const GramPlan g = gram_plan(N, k); return N * g.JS * g.tiles; }
int gram_js(int N, int k) { return gram_plan(N, k).JS; }

static size_t gram_smem(int Dx, int Dz, bool pass2) {
  const size_t img = gram_dims(Dx, Dz).img_bytes;
  return img + 2 * (img + kNormBytes + (pass2 ? (size_t)kXTBytes : 0)) + (size_t)kGT * 8 + 2 * kGT * 8 + 32 + 64 + 64;
}

// mean -> scale -> tile images + norms; clears the reduction tickets.  ws: gram_ws_bytes()
int gram_prepare(const float* x, const float* z, int N, int k, int Dx, int Dz, void* ws, cudaStream_t s) {
  const GramWs w = gram_ws_layout(N, k, Dx, Dz);
  char* b = reinterpret_cast<char*>(ws);
  const int tiles = (k + kGT - 1) / kGT;
  ndiv_prepack_kernel<<<dim3(4 * tiles, N), 256, 0, s>>>(x, z, k, Dx, Dz, reinterpret_cast<uint8_t*>(b + w.img),
                                                      reinterpret_cast<uint8_t*>(b + w.imgT),
                                                      reinterpret_cast<float*>(b + w.mu),
                                                      reinterpret_cast<float2*>(b + w.norms),
                                                      reinterpret_cast<unsigned*>(b + w.tickets), 2 * N * tiles + 2);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

static void gram_fill(GramParams& p, const float* x, const float* z, int N, int k, int Dx, int Dz, void* ws, int row0,
                      int rows, const GramPlan* plan = nullptr) {
  const GramWs w = gram_ws_layout(N, k, Dx, Dz);
  const GramPlan g = plan ? *plan : gram_plan(N, k);
  char* b = reinterpret_cast<char*>(ws);
  p.img = reinterpret_cast<const uint8_t*>(b + w.img);
  p.norms = reinterpret_cast<const float2*>(b + w.norms);
  p.imgT = reinterpret_cast<const uint8_t*>(b + w.imgT);
  p.mu = reinterpret_cast<const float*>(b + w.mu);
  p.tickets = reinterpret_cast<unsigned*>(b + w.tickets);
  p.x = x; p.z = z;
  p.k = k; p.Dx = Dx; p.Dz = Dz; p.tiles = g.tiles; p.tps = g.tps; p.JS = g.JS; p.row0 = row0; p.rows = rows;
}

static int gram_set_smem(const void* fn, size_t bytes) {
  int dev = 0;
  NDP_CUDA(cudaGetDevice(&dev));
  NDP_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return NDP_OK;
}

// pass 1 on tensor cores: final row sums sx/sz [N][k] for rows [row0,row0+rows); psx/psz [JS][N][k] scratch (JS > 1)
int gram_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows, void* ws, float* sx,
                 float* sz, float* psx, float* psz, cudaStream_t s) {
  GramParams p{};
  gram_fill(p, x, z, N, k, Dx, Dz, ws, row0, rows);
  p.psx = psx; p.psz = psz; p.out_sx = sx; p.out_sz = sz;
  const size_t sm = gram_smem(Dx, Dz, false);
  if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<false>, sm)) return rc;
  ndiv_gram_kernel<false><<<dim3((rows + kGT - 1) / kGT, p.JS, N), kGNT, sm, s>>>(p);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// pass 2 on tensor cores: final loss[0] and gradient grad_x (nullable) for rows [row0,row0+rows);
// loss_part [gram_loss_parts] and gpart [JS][N][k][Dx] (JS > 1, with grad) are scratch
int gram_loss_grad(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx, int Dz,
                   float margin, int row0, int rows, void* ws, float* loss_part, float* gpart, float* loss, float* grad_x,
                   cudaStream_t s) {
  GramParams p{};
  gram_fill(p, x, z, N, k, Dx, Dz, ws, row0, rows);
  p.sx = sx; p.sz = sz; p.margin = margin; p.loss_part = loss_part; p.gpart = gpart; p.loss = loss; p.grad = grad_x;
  const size_t sm = gram_smem(Dx, Dz, true);
  if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<true>, sm)) return rc;
  ndiv_gram_kernel<true><<<dim3((rows + kGT - 1) / kGT, p.JS, N), kGNT, sm, s>>>(p);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---------------------------------------------------------------- row-sharded forward + backward, one group (N = 1)
// SURVEY 8e row 3 as ONE stream-ordered call per rank: pre-pack once, pass 1 on the rank's 128-aligned row block,
// in-place all-gather of the [sx | sz] blocks, pass 2, in-place all-gather of [grad rows | loss partial]; the kernels
// write straight into the rank's slot of the exchange buffers (pointer offsets), two tiny kernels unpack.  The column
// split adapts to the rows of the CALL (a rank with 1/P of the rows still fills the GPU), so results equal the
// single-GPU call to fp32 round-off of the split order, not bit for bit; they are identical on every rank.
constexpr int kShardMaxJS = 32;
static GramPlan gram_plan_rows(int k, int row_tiles) {
  GramPlan g;
  g.tiles = (k + kGT - 1) / kGT;
  long long want = (148 * 2) / (long long)(row_tiles > 0 ? row_tiles : 1);
  if (want < 1) want = 1;
  if (want > kShardMaxJS) want = kShardMaxJS;            // bounds the partial buffers
  if (want > g.tiles) want = g.tiles;
  g.tps = (int)((g.tiles + want - 1) / want);
  g.JS = (g.tiles + g.tps - 1) / g.tps;
  return g;
}
struct ShardWs { size_t gram, part, psum, st_s, st_g, total; int B; };
static ShardWs shard_ws_layout(int k, int Dx, int Dz, int world) {
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  ShardWs w;
  w.B = ((k + world - 1) / world + kGT - 1) / kGT * kGT;
  const size_t tiles = (size_t)(k + kGT - 1) / kGT;
  w.gram = 0;
  w.part = al(gram_ws_bytes(1, k, Dx, Dz));
  w.psum = w.part + al(kShardMaxJS * tiles * sizeof(float));
  const size_t ps = (size_t)2 * kShardMaxJS * k * sizeof(float), gp = (size_t)kShardMaxJS * k * Dx * sizeof(float);
  w.st_s = w.psum + al(ps > gp ? ps : gp);
  w.st_g = w.st_s + al((size_t)world * 2 * w.B * sizeof(float));
  w.total = w.st_g + al((size_t)world * ((size_t)w.B * Dx + 1) * sizeof(float));
  return w;
}
__global__ void ndiv_unpack_s_kernel(const float* __restrict__ st, int B, int k, float* __restrict__ sx, float* __restrict__ sz) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  const float* blk = st + (size_t)(j / B) * 2 * B;
  sx[j] = blk[j % B];
  sz[j] = blk[B + j % B];
}
__global__ void ndiv_unpack_g_kernel(const float* __restrict__ st, int P, int B, int k, int Dx, float* __restrict__ grad,
                                     float* __restrict__ loss) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t per = (size_t)B * Dx + 1;
  if (grad && e < (size_t)k * Dx) {
    const size_t j = e / Dx;
    grad[e] = st[(j / B) * per + (j % B) * Dx + e % Dx];
  }
  if (e == 0) {
    float t = 0.f;
    for (int r = 0; r < P; ++r) t += st[(size_t)r * per + per - 1];     // rank order: identical bits on every rank
    loss[0] = t;
  }
}
// ---- the same exchange over the peer window (peer.cuh): the pass kernels' last CTA publishes the rank's slot, these
// kernels wait for every rank's flag and PULL the slots over NVLink (no staging copy, no library collective)
struct ShardWin { size_t s_off, g_off, half, need; };
static ShardWin shard_win_layout(int B, int Dx, size_t win_bytes) {
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  ShardWin w;
  const size_t per_parity = al((size_t)2 * B * 4) + al(((size_t)B * Dx + 1) * 4);
  w.need = 2 * per_parity;
  // the ndiv region is the first half of the window's CAPACITY (the second half holds plan records, k3_select.cu);
  // its parity halves split the capacity (identical on every rank), not this shape's need, so calls of different
  // shapes on one communicator never overlap a region a slower peer may still be pulling from
  w.half = win_bytes > kWinHeader ? ((win_bytes - kWinHeader) / 4) / 256 * 256 : 0;
  w.s_off = 0;
  w.g_off = al((size_t)2 * B * 4);
  return w;
}
__global__ void peer_signal_kernel(PeerWin w, int slot, unsigned epoch) { peer_signal_all(w, slot, epoch); }
__global__ void ndiv_pull_s_kernel(PeerWin w, size_t off, unsigned epoch, int B, int k, float* __restrict__ sx,
                                   float* __restrict__ sz) {
  if ((int)threadIdx.x < w.world) peer_wait(w, 0, threadIdx.x, epoch);
  __syncthreads();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= k) return;
  const float* blk = reinterpret_cast<const float*>(w.peer[j / B] + off);
  sx[j] = ld_peer(blk + j % B);
  sz[j] = ld_peer(blk + B + j % B);
}
__global__ void ndiv_pull_g_kernel(PeerWin w, size_t off, unsigned epoch, int B, int k, int Dx, float* __restrict__ grad,
                                   float* __restrict__ loss) {
  if ((int)threadIdx.x < w.world) peer_wait(w, 1, threadIdx.x, epoch);
  __syncthreads();
  const size_t e = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;       // Dx % 4 == 0
  if (grad && e < (size_t)k * Dx) {
    const size_t j = e / Dx;
    const float* slot = reinterpret_cast<const float*>(w.peer[j / B] + off);
    *reinterpret_cast<float4*>(grad + e) = ld_peer4(reinterpret_cast<const float4*>(slot + (j % B) * Dx + e % Dx));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    float t = 0.f;
    for (int r = 0; r < w.world; ++r)                       // rank order: identical bits on every rank
      t += ld_peer(reinterpret_cast<const float*>(w.peer[r] + off) + (size_t)B * Dx);
    loss[0] = t;
  }
}
}  // namespace ndp

using namespace ndp;
#ifdef NDP_TC_PROF
extern "C" int ndp_debug_gram_prof(unsigned long long* out16, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out16, g_gram_prof, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {}; cudaMemcpyToSymbol(g_gram_prof, z, sizeof(z)); }
  return NDP_OK;
}
#endif
extern "C" int ndp_ndiv_sharded_workspace_bytes(int k, int Dx, int Dz, int world, size_t* bytes) {
  NDP_CHECK_ARG(bytes && k >= 1 && Dx >= 1 && Dz >= 1 && world >= 1, "ndiv_sharded: bad arguments");
  *bytes = shard_ws_layout(k, Dx, Dz, world).total;
  return NDP_OK;
}
extern "C" int ndp_ndiv_sharded_forward_backward(ndp_comm* c, const float* x, const float* z, int k, int Dx, int Dz,
                                                 float margin, float* sx, float* sz, float* loss, float* grad_x, void* ws,
                                                 size_t ws_bytes, void* stream) {
  NDP_CHECK_ARG(c && x && z && sx && sz && loss, "ndiv_sharded: null argument");
  if (int rc = device_check()) return rc;
  NDP_CHECK_ARG(gram_eligible(1, k, Dx, Dz, 0), "ndiv_sharded: one group with k >= 1536, Dx,Dz <= 32, Dx %% 4 == 0 (use ShardedNdiv's generic path otherwise)");
  const int world = ndp_comm_world(c), rank = ndp_comm_rank(c);
  const ShardWs w = shard_ws_layout(k, Dx, Dz, world);
  if (!ws || ws_bytes < w.total) { set_error("ndiv_sharded: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  NDP_CHECK_ARG(((uintptr_t)ws & 255) == 0, "ndiv_sharded: workspace must be 256-byte aligned");
  NDP_CHECK_ARG(!grad_x || ((uintptr_t)grad_x & 15) == 0, "ndiv_sharded: grad_x must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  char* b = reinterpret_cast<char*>(ws);
  const int B = w.B;
  const int row0 = rank * B < k ? rank * B : k;
  const int rows = (row0 + B < k ? row0 + B : k) - row0;
  float* st_s = reinterpret_cast<float*>(b + w.st_s);
  float* st_g = reinterpret_cast<float*>(b + w.st_g);
  const size_t per = (size_t)B * Dx + 1;
  const GramPlan plan = gram_plan_rows(k, (rows + kGT - 1) / kGT);
  float* psum = reinterpret_cast<float*>(b + w.psum);
  // ---- exchange over the peer window when the ranks can map each other (one node, NVLink) and no graph is being
  // captured (the epoch is a launch argument); otherwise the NCCL all-gathers below
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  NDP_CUDA(cudaStreamIsCapturing(s, &cap));
  PeerWin pw{};
  if (cap == cudaStreamCaptureStatusNone) {
    if (int rc = comm_window(c, 2 * shard_win_layout(B, Dx, 0).need + 4096, s, &pw)) return rc;
  }
  if (pw.world > 0) {
    const ShardWin wl = shard_win_layout(B, Dx, c->win_bytes);
    const unsigned epoch = ++c->epoch;
    const size_t base = kWinHeader + (size_t)(epoch & 1u) * wl.half;
    float* my_s = reinterpret_cast<float*>(c->win + base + wl.s_off);
    float* my_g = reinterpret_cast<float*>(c->win + base + wl.g_off);
    if (int rc = gram_prepare(x, z, 1, k, Dx, Dz, b + w.gram, s)) return rc;
    if (rows > 0) {
      GramParams p{};
      gram_fill(p, x, z, 1, k, Dx, Dz, b + w.gram, row0, rows, &plan);
      p.psx = psum; p.psz = psum + (size_t)plan.JS * k;
      p.out_sx = my_s - row0;                                 // row i lands at slot[i - row0]
      p.out_sz = my_s + B - row0;
      p.pw = pw; p.sig_slot = 0; p.sig_epoch = epoch;
      const size_t sm = gram_smem(Dx, Dz, false);
      if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<false>, sm)) return rc;
      ndiv_gram_kernel<false><<<dim3((rows + kGT - 1) / kGT, plan.JS, 1), kGNT, sm, s>>>(p);
    } else {
      peer_signal_kernel<<<1, 32, 0, s>>>(pw, 0, epoch);
    }
    NDP_LAUNCH_CHECK();
    ndiv_pull_s_kernel<<<(k + 255) / 256, 256, 0, s>>>(pw, base + wl.s_off, epoch, B, k, sx, sz);
    NDP_LAUNCH_CHECK();
    if (rows > 0) {
      GramParams p{};
      gram_fill(p, x, z, 1, k, Dx, Dz, b + w.gram, row0, rows, &plan);
      p.sx = sx; p.sz = sz; p.margin = margin;
      p.loss_part = reinterpret_cast<float*>(b + w.part);
      p.gpart = psum;
      p.loss = my_g + (size_t)B * Dx;
      p.grad = grad_x ? my_g - (size_t)row0 * Dx : nullptr;
      p.pw = pw; p.sig_slot = 1; p.sig_epoch = epoch;
      const size_t sm = gram_smem(Dx, Dz, true);
      if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<true>, sm)) return rc;
      ndiv_gram_kernel<true><<<dim3((rows + kGT - 1) / kGT, plan.JS, 1), kGNT, sm, s>>>(p);
    } else {
      NDP_CUDA(cudaMemsetAsync(my_g + (size_t)B * Dx, 0, sizeof(float), s));
      peer_signal_kernel<<<1, 32, 0, s>>>(pw, 1, epoch);
    }
    NDP_LAUNCH_CHECK();
    const size_t n4 = grad_x ? (size_t)k * Dx / 4 : 1;
    ndiv_pull_g_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, s>>>(pw, base + wl.g_off, epoch, B, k, Dx, grad_x, loss);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  if (int rc = gram_prepare(x, z, 1, k, Dx, Dz, b + w.gram, s)) return rc;
  if (rows > 0) {
    GramParams p{};
    gram_fill(p, x, z, 1, k, Dx, Dz, b + w.gram, row0, rows, &plan);
    p.psx = psum; p.psz = psum + (size_t)plan.JS * k;
    p.out_sx = st_s + (size_t)rank * 2 * B - row0;               // row i lands at slot[rank][0][i - row0]
    p.out_sz = st_s + (size_t)rank * 2 * B + B - row0;
    const size_t sm = gram_smem(Dx, Dz, false);
    if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<false>, sm)) return rc;
    ndiv_gram_kernel<false><<<dim3((rows + kGT - 1) / kGT, plan.JS, 1), kGNT, sm, s>>>(p);
    NDP_LAUNCH_CHECK();
  }
  if (int rc = ndp_comm_allgather_f32(c, st_s + (size_t)rank * 2 * B, st_s, 2 * B, stream)) return rc;
  ndiv_unpack_s_kernel<<<(k + 255) / 256, 256, 0, s>>>(st_s, B, k, sx, sz);
  NDP_LAUNCH_CHECK();
  if (rows > 0) {
    GramParams p{};
    gram_fill(p, x, z, 1, k, Dx, Dz, b + w.gram, row0, rows, &plan);
    p.sx = sx; p.sz = sz; p.margin = margin;
    p.loss_part = reinterpret_cast<float*>(b + w.part);
    p.gpart = psum;
    p.loss = st_g + (size_t)rank * per + per - 1;
    p.grad = grad_x ? st_g + (size_t)rank * per - (size_t)row0 * Dx : nullptr;   // row i at slot[rank][(i - row0) * Dx]
    const size_t sm = gram_smem(Dx, Dz, true);
    if (int rc = gram_set_smem((const void*)ndiv_gram_kernel<true>, sm)) return rc;
    ndiv_gram_kernel<true><<<dim3((rows + kGT - 1) / kGT, plan.JS, 1), kGNT, sm, s>>>(p);
    NDP_LAUNCH_CHECK();
  } else {
    NDP_CUDA(cudaMemsetAsync(st_g + (size_t)rank * per + per - 1, 0, sizeof(float), s));
  }
  if (int rc = ndp_comm_allgather_f32(c, st_g + (size_t)rank * per, st_g, (int)per, stream)) return rc;
  const size_t n = grad_x ? (size_t)k * Dx : 1;
  ndiv_unpack_g_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(st_g, world, B, k, Dx, grad_x, loss);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
