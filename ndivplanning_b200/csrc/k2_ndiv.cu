// K2 — normalized diversification loss (reference diversity.py:8-41), forward + backward,
// without ever materialising a k x k matrix in HBM.
//
//   d_ij  = || v_i - v_j ||_2                      compute_pairwise        diversity.py:8-9
//   delta = d_ij / sum_j d_ij (normaliser detached) compute_pair_distance   diversity.py:12-19
//   L     = sum relu(0.8 * delta^z_ij - delta^x_ij) compute_pairwise_divergence :36-41
//   dL/dx_i = sum_j What_ij (x_i - x_j),  What_ij = -(M_ij/sx_i + M_ji/sx_j) / dx_ij  (0 if dx_ij==0)
//
// Arithmetic is the reference's own (direct differences, fp32, IEEE sqrt) so parity does not depend on
// a Gram-form cancellation: the loss exists to separate NEAR-DUPLICATE outputs, exactly the regime where
// |a|^2+|b|^2-2ab loses all digits.  Two passes over the pair tiles:
//   pass 1  row sums sx, sz                          (ndp_ndiv_rowsums)
//   pass 2  loss partials + What tile + What*X GEMM  (ndp_ndiv_loss_grad)
// 64 x 128 pair tiles, 4x4 register sub-tiles, every feature of a tile staged once in shared memory
// (transposed, conflict-free), gradient product register-tiled 4 rows x 8 dims x 32 j per thread.
// Algorithmic HBM bytes: 4*N*k*(Dz+2*Dx) + 8*N*k (nothing k x k ever leaves the SM).
#include "common.cuh"

namespace ndp {

// Tiling: a block owns kTI rows i and sweeps a contiguous range of kTJ-wide column tiles j (one of JS
// "j-splits" of the row, so that k/kTI * JS * N blocks fill the 148 SMs); partial row sums / gradients of
// the splits are added in fixed order by a small reduce kernel -> deterministic, no atomics, and a row's
// result does not depend on how the caller blocks the rows (multi-GPU row sharding is bit-identical).
static constexpr int kTI = 64;              // rows per block
static constexpr int kTJ = 128;             // columns per tile
static constexpr int kNT = 256;             // threads per block
static constexpr int kDC = 16;              // feature dims per chunk (x and z are padded to chunks separately)
static constexpr int kSG = 3;               // chunks staged together (48 dims: x32 | z16 in one round)
static constexpr int kSD = kSG * kDC;
static constexpr int kSI = 2 * kTI + 8;     // i rows are staged DUPLICATED ({a,a} pairs feed the packed f32x2 ops): 136
                                            // smem row strides in floats, 16-byte aligned
static constexpr int kSJ = kTJ + 4;
static constexpr int kGD = 32;              // gradient dims per GEMM pass
static constexpr int kXJ = kGD + 4;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct NdivDims {
  int cx, cz;        // 16-dim chunks of x / z
  int groups;        // staging rounds of kSG chunks over the concatenated [x | z] chunk list
};
__host__ __device__ inline NdivDims ndiv_dims(int Dx, int Dz) {
  NdivDims d;
  d.cx = (Dx + kDC - 1) / kDC;
  d.cz = (Dz + kDC - 1) / kDC;
  d.groups = (d.cx + d.cz + kSG - 1) / kSG;
  return d;
}

// Stage chunks [c0, c0+nc) of the concatenated chunk list of rows [r0, r0+R) transposed into s[dim][row]
// (zero padded).  Warp item = 4 rows x 8 dims: 4 x 32-byte global segments in, bank-conflict-free smem out.
template <int R, int STRIDE, bool DUP>
__device__ __forceinline__ void stage_rows(float* __restrict__ s, const float* __restrict__ x, const float* __restrict__ z,
                                           int k, int Dx, int Dz, int cx, int r0, int c0, int nc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rr = lane & 3, dd = lane >> 2;
  for (int rq = warp; rq < R / 4; rq += kNT / 32) {   // row quad
    const int r = r0 + rq * 4 + rr;
    const bool rv = r < k;
    const float* xr = x + (size_t)r * Dx;
    const float* zr = z + (size_t)r * Dz;
    float* dst = s + dd * STRIDE + (rq * 4 + rr) * (DUP ? 2 : 1);
    for (int c = 0; c < nc; ++c) {
      const bool isx = (c0 + c) < cx;
      const float* src = isx ? xr : zr;
      const int D = isx ? Dx : Dz;
      const int d = (isx ? (c0 + c) : (c0 + c - cx)) * kDC + dd;
      const float v0 = (rv && d < D) ? __ldg(src + d) : 0.f;
      const float v1 = (rv && d + 8 < D) ? __ldg(src + d + 8) : 0.f;
      if (DUP) {
        *reinterpret_cast<float2*>(dst + (c * kDC) * STRIDE) = make_float2(v0, v0);
        *reinterpret_cast<float2*>(dst + (c * kDC + 8) * STRIDE) = make_float2(v1, v1);
      } else {
        dst[(c * kDC) * STRIDE] = v0;
        dst[(c * kDC + 8) * STRIDE] = v1;
      }
    }
  }
}

// squared distances of the 4x4 pair sub-tile (rows rb..rb+3, cols cb..cb+3) over one staged chunk.  A warp covers
// 16 rows x 32 cols (lane = 4 row quads x 8 col quads): both operand loads are a single 64/128-byte wavefront.
typedef unsigned long long u64;   // a packed pair of fp32 (sm_100 f32x2 arithmetic: FADD2 / FFMA2 in SASS)
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ u64 dup2(float v) { u64 r; asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(v)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }

// acc[p][h] = packed squared distances of pairs (row rb+p, cols cb+2h, cb+2h+1)
__device__ __forceinline__ void sqdist_chunk(u64 (&acc)[4][2], const float* __restrict__ si, const float* __restrict__ sj,
                                             int rb, int cb) {
#pragma unroll
  for (int d = 0; d < kDC; ++d) {
    const ulonglong2 a01 = *reinterpret_cast<const ulonglong2*>(si + d * kSI + rb * 2);       // {a0,a0} {a1,a1}
    const ulonglong2 a23 = *reinterpret_cast<const ulonglong2*>(si + d * kSI + rb * 2 + 4);   // {a2,a2} {a3,a3}
    const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(sj + d * kSJ + cb);            // {b0,b1} {b2,b3}
    const u64 a[4] = {a01.x, a01.y, a23.x, a23.y}, b[2] = {b4.x, b4.y};
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const u64 df = sub2(a[p], b[h]);
        acc[p][h] = fma2(df, df, acc[p][h]);
      }
  }
}

// all chunks of one matrix kind present in the staged group: chunks [c0,c0+nc) of the list, x-type iff c < cx
__device__ __forceinline__ void sqdist_group(u64 (&ax)[4][2], u64 (&az)[4][2], const float* si, const float* sj,
                                             int cx, int c0, int nc, int rb, int cb) {
  for (int c = 0; c < nc; ++c) {
    if (c0 + c < cx) sqdist_chunk(ax, si + c * kDC * kSI, sj + c * kDC * kSJ, rb, cb);
    else             sqdist_chunk(az, si + c * kDC * kSI, sj + c * kDC * kSJ, rb, cb);
  }
}

__device__ __forceinline__ void zero42(u64 (&a)[4][2]) {
#pragma unroll
  for (int p = 0; p < 4; ++p) a[p][0] = a[p][1] = 0ull;
}
__device__ __forceinline__ void unpack44(const u64 (&a)[4][2], float (&f)[4][4]) {
#pragma unroll
  for (int p = 0; p < 4; ++p) { unpack2(a[p][0], f[p][0], f[p][1]); unpack2(a[p][1], f[p][2], f[p][3]); }
}

// ---------------------------------------------------------------- pass 1: row sums
// grid (row blocks, JS, N).  out: psx/psz [JS][N][k] partial sums (JS == 1: the sums themselves)
__global__ void __launch_bounds__(kNT, 3)
ndiv_rowsum_kernel(const float* __restrict__ x, const float* __restrict__ z, int k, int Dx, int Dz,
                   int row0, int rows, int tps, float* __restrict__ psx, float* __restrict__ psz,
                   const int* __restrict__ run_if) {     // nullable: the whole grid exits when *run_if == 0
  if (run_if && *run_if == 0) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* si = reinterpret_cast<float*>(smem_raw);
  float* sj = si + kSD * kSI;
  float (*srs)[2][kTI] = reinterpret_cast<float (*)[2][kTI]>(sj + kSD * kSJ);
  const int n = blockIdx.z, js = blockIdx.y, N = gridDim.z;
  const int i0 = row0 + blockIdx.x * kTI;
  const int iend = (row0 + rows) < k ? (row0 + rows) : k;
  const int ntiles = (k + kTJ - 1) / kTJ;
  const int jt0 = js * tps, jt1 = (jt0 + tps) < ntiles ? (jt0 + tps) : ntiles;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rb = (warp & 3) * 16 + (lane >> 3) * 4;    // warp = 16 rows x 32 cols, lane = (row quad, col quad)
  const int wc = warp >> 2, lc = lane & 7;
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  const NdivDims nd = ndiv_dims(Dx, Dz);
  const int nchunks = nd.cx + nd.cz;
  const bool hoist = nd.groups == 1;                  // all dims fit one staging round: stage the i rows once
  if (hoist) stage_rows<kTI, kSI, true>(si, xg, zg, k, Dx, Dz, nd.cx, i0, 0, nchunks);
  float rsx[4] = {0.f, 0.f, 0.f, 0.f}, rsz[4] = {0.f, 0.f, 0.f, 0.f};
  for (int jt = jt0; jt < jt1; ++jt) {
    const int j0 = jt * kTJ;
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      const int cb = (wc * 2 + half) * 32 + lc * 4;
      u64 ax2[4][2], az2[4][2];
      zero42(ax2); zero42(az2);
      for (int g = 0; g < nd.groups; ++g) {
        const int c0 = g * kSG, nc = (nchunks - c0) < kSG ? (nchunks - c0) : kSG;
        if (half == 0 || !hoist) {                    // hoisted case: the j rows are staged once per tile
          __syncthreads();
          if (!hoist) stage_rows<kTI, kSI, true>(si, xg, zg, k, Dx, Dz, nd.cx, i0, c0, nc);
          stage_rows<kTJ, kSJ, false>(sj, xg, zg, k, Dx, Dz, nd.cx, j0, c0, nc);
          __syncthreads();
        }
        sqdist_group(ax2, az2, si, sj, nd.cx, c0, nc, rb, cb);
      }
      float ax[4][4], az[4][4];
      unpack44(ax2, ax); unpack44(az2, az);
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (j0 + cb + q < k) {
            rsx[p] += sqrtf(ax[p][q]);
            rsz[p] += sqrtf(az[p][q]);
          }
    }
  }
  // row sums: 8 column-quad lanes (shuffles), then the two column halves of the block (shared memory), fixed order
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    float vx = rsx[p], vz = rsz[p];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      vx += __shfl_xor_sync(0xffffffffu, vx, o);
      vz += __shfl_xor_sync(0xffffffffu, vz, o);
    }
    if (lc == 0) { srs[0][wc][rb + p] = vx; srs[1][wc][rb + p] = vz; }
  }
  __syncthreads();
  if (threadIdx.x < kTI) {
    const int i = i0 + threadIdx.x;
    if (i < iend) {
      const size_t o = ((size_t)js * N + n) * k + i;
      psx[o] = srs[0][0][threadIdx.x] + srs[0][1][threadIdx.x];
      psz[o] = srs[1][0][threadIdx.x] + srs[1][1][threadIdx.x];
    }
  }
}

// out[n][i] = sum_js part[js][n][i] in fixed order, rows [row0,row0+rows)
__global__ void ndiv_rowsum_reduce_kernel(const float* __restrict__ psx, const float* __restrict__ psz, int N, int k,
                                          int JS, int row0, int rows, float* __restrict__ sx, float* __restrict__ sz,
                                          const int* __restrict__ run_if) {
  if (run_if && *run_if == 0) return;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= N * rows) return;
  const int n = t / rows, i = row0 + t % rows;
  float a = 0.f, b = 0.f;
  for (int js = 0; js < JS; ++js) {
    const size_t o = ((size_t)js * N + n) * k + i;
    a += psx[o];
    b += psz[o];
  }
  sx[(size_t)n * k + i] = a;
  sz[(size_t)n * k + i] = b;
}

// ---------------------------------------------------------------- pass 2: loss + gradient
struct NdivSmem2 {
  float si[kSD * kSI];
  float sj[kSD * kSJ];
  float sW[kTI * kSJ];          // What tile; reused for the final cross-warp gradient reduction
  float sxj[kTJ * kXJ];         // x_j rows of the current gradient dim pass, j-major
  float2 ri[kTI];               // {1/sx_i, 1/sz_i}
  float2 rj[kTJ];               // {1/sx_j, 1/sz_j}
  float red[kNT / 32];
  float sw[4 * kTI];
};

// epilogue of a 4x4 pair sub-tile: loss terms, What -> sW.  EDGE: tile touches the row/column boundary.
template <bool EDGE, bool WITH_GRAD>
__device__ __forceinline__ void ndiv_epilogue(const float (&ax)[4][4], const float (&az)[4][4], NdivSmem2& sm, float margin,
                                              int rb, int cb, int i_valid, int j_valid, float& lsum) {
  float2 cj[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) cj[q] = sm.rj[cb + q];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const float2 ri = sm.ri[rb + p];
    float w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float ddx = sqrtf(ax[p][q]), ddz = sqrtf(az[p][q]);
      // relu(margin * dz/sz_i - dx/sx_i), diversity.py:18,40 (normalisers as reciprocals)
      const float hij = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, ri.y)), __fmul_rn(ddx, ri.x));
      const float hji = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, cj[q].y)), __fmul_rn(ddx, cj[q].x));
      const bool ok = !EDGE || ((rb + p) < i_valid && (cb + q) < j_valid);
      if (ok) lsum += !(hij <= 0.f) ? hij : 0.f;       // NaN propagates (k == 1 -> 0/0, like the reference)
      if (WITH_GRAD) {
        const float wij = (hij > 0.f ? ri.x : 0.f) + (hji > 0.f ? cj[q].x : 0.f);
        w[q] = (ok && ddx > 0.f) ? __fdividef(-wij, ddx) : 0.f;   // torch.norm subgradient: 0 at d == 0
      }
    }
    if (WITH_GRAD) *reinterpret_cast<float4*>(sm.sW + (rb + p) * kSJ + cb) = make_float4(w[0], w[1], w[2], w[3]);
  }
}

// The gradient product x_i * sum_j What_ij - sum_j What_ij x_j is evaluated on rows centred on the group's first row
// (x - x_0): the two terms cancel to O(|x_i - x_j|), so without centring a tight cluster far from the origin (the S
// samples of one code during GAN training) would lose |x| / |x_i - x_j| in relative accuracy.
__device__ __forceinline__ float4 centre4(const float* __restrict__ xg, int d, int Dx) {
  float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
  if (d < Dx) c.x = __ldg(xg + d);
  if (d + 1 < Dx) c.y = __ldg(xg + d + 1);
  if (d + 2 < Dx) c.z = __ldg(xg + d + 2);
  if (d + 3 < Dx) c.w = __ldg(xg + d + 3);
  return c;
}

// grid (row blocks, JS, N).  gpart [JS][N][k][Dx] (JS == 1: grad_x itself), loss_part [gridDim.x*JS*N].
// NDP = number of 32-dim gradient passes held in registers; dims [dpass0, dpass0 + 32*NDP) are produced.
template <int NDP, bool WITH_GRAD>
__global__ void __launch_bounds__(kNT, NDP == 1 ? 2 : 1)
ndiv_loss_grad_kernel(const float* __restrict__ x, const float* __restrict__ z, const float* __restrict__ sx,
                      const float* __restrict__ sz, int k, int Dx, int Dz, float margin, int row0, int rows,
                      int tps, int dpass0, float* __restrict__ loss_part, float* __restrict__ gpart,
                      const int* __restrict__ run_if) {
  if (run_if && *run_if == 0) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  NdivSmem2& sm = *reinterpret_cast<NdivSmem2*>(smem_raw);
  const int n = blockIdx.z, js = blockIdx.y, N = gridDim.z;
  const int i0 = row0 + blockIdx.x * kTI;
  const int iend = (row0 + rows) < k ? (row0 + rows) : k;
  const int ntiles = (k + kTJ - 1) / kTJ;
  const int jt0 = js * tps, jt1 = (jt0 + tps) < ntiles ? (jt0 + tps) : ntiles;
  const int tid = threadIdx.x, ty = tid >> 5, tx = tid & 31;
  const int rb = (ty & 3) * 16 + (tx >> 3) * 4;        // distance phase: warp = 16 rows x 32 cols, lane = (row quad, col quad)
  const int wc = ty >> 2, lc = tx & 7;
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  const float* sxg = sx + (size_t)n * k;
  const float* szg = sz + (size_t)n * k;
  const NdivDims nd = ndiv_dims(Dx, Dz);
  const int nchunks = nd.cx + nd.cz;
  const bool hoist = nd.groups == 1;
  // GEMM roles: warp = (j quarter, row half); lane = (row group, 8-dim group)
  const int jq = ty & 3, rg16 = (tx >> 2) + 8 * (ty >> 2), dg = tx & 3;

  if (tid < kTI) {
    const int i = i0 + tid;
    sm.ri[tid] = i < k ? make_float2(1.f / __ldg(sxg + i), 1.f / __ldg(szg + i)) : make_float2(0.f, 0.f);
  }
  if (hoist) stage_rows<kTI, kSI, true>(sm.si, xg, zg, k, Dx, Dz, nd.cx, i0, 0, nchunks);
  float gacc[NDP][4][8];
  float swp[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int e = 0; e < NDP; ++e)
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int d = 0; d < 8; ++d) gacc[e][a][d] = 0.f;
  float lsum = 0.f;

  for (int jt = jt0; jt < jt1; ++jt) {
    const int j0 = jt * kTJ;
    const bool edge = (i0 + kTI > iend) || (j0 + kTJ > k);
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      const int cb = (wc * 2 + half) * 32 + lc * 4;
      u64 ax2[4][2], az2[4][2];
      zero42(ax2); zero42(az2);
      for (int g = 0; g < nd.groups; ++g) {
        const int c0 = g * kSG, nc = (nchunks - c0) < kSG ? (nchunks - c0) : kSG;
        if (half == 0 || !hoist) {
          __syncthreads();                             // previous tile's GEMM / previous group's readers are done
          if (!hoist) stage_rows<kTI, kSI, true>(sm.si, xg, zg, k, Dx, Dz, nd.cx, i0, c0, nc);
          stage_rows<kTJ, kSJ, false>(sm.sj, xg, zg, k, Dx, Dz, nd.cx, j0, c0, nc);
          if (half == 0 && g == 0) {
            if (tid < kTJ) {
              const int j = j0 + tid;
              sm.rj[tid] = j < k ? make_float2(1.f / __ldg(sxg + j), 1.f / __ldg(szg + j)) : make_float2(0.f, 0.f);
            }
            if (WITH_GRAD) {                           // x_j rows of gradient pass 0, j-major, float4 granules
              for (int t = tid; t < kTJ * (kGD / 4); t += kNT) {
                const int j = t >> 3, q = t & 7, d = dpass0 + q * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (j0 + j < k) {
                  const float* src = xg + (size_t)(j0 + j) * Dx + d;
                  if ((Dx & 3) == 0 && d + 3 < Dx) v = __ldg(reinterpret_cast<const float4*>(src));
                  else {
                    if (d < Dx) v.x = __ldg(src);
                    if (d + 1 < Dx) v.y = __ldg(src + 1);
                    if (d + 2 < Dx) v.z = __ldg(src + 2);
                    if (d + 3 < Dx) v.w = __ldg(src + 3);
                  }
                  const float4 c = centre4(xg, d, Dx);
                  v.x -= c.x; v.y -= c.y; v.z -= c.z; v.w -= c.w;
                }
                *reinterpret_cast<float4*>(sm.sxj + j * kXJ + q * 4) = v;
              }
            }
          }
          __syncthreads();
        }
        sqdist_group(ax2, az2, sm.si, sm.sj, nd.cx, c0, nc, rb, cb);
      }
      float ax[4][4], az[4][4];
      unpack44(ax2, ax); unpack44(az2, az);
      if (edge) ndiv_epilogue<true, WITH_GRAD>(ax, az, sm, margin, rb, cb, iend - i0, k - j0, lsum);
      else      ndiv_epilogue<false, WITH_GRAD>(ax, az, sm, margin, rb, cb, kTI, kTJ, lsum);
    }
    if (WITH_GRAD) {
      // grad_i -= sum_j What_ij x_j : [64 x 128] x [128 x 32] per pass, 4 rows x 8 dims x 32 j per thread
#pragma unroll
      for (int e = 0; e < NDP; ++e) {
        if (e > 0) {
          if (dpass0 + e * kGD >= Dx) break;
          __syncthreads();                             // previous pass has consumed sxj
          for (int t = tid; t < kTJ * (kGD / 4); t += kNT) {
            const int j = t >> 3, q = t & 7, d = dpass0 + e * kGD + q * 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j0 + j < k) {
              const float* src = xg + (size_t)(j0 + j) * Dx + d;
              if ((Dx & 3) == 0 && d + 3 < Dx) v = __ldg(reinterpret_cast<const float4*>(src));
              else {
                if (d < Dx) v.x = __ldg(src);
                if (d + 1 < Dx) v.y = __ldg(src + 1);
                if (d + 2 < Dx) v.z = __ldg(src + 2);
                if (d + 3 < Dx) v.w = __ldg(src + 3);
              }
              const float4 c = centre4(xg, d, Dx);
              v.x -= c.x; v.y -= c.y; v.z -= c.z; v.w -= c.w;
            }
            *reinterpret_cast<float4*>(sm.sxj + j * kXJ + q * 4) = v;
          }
        }
        __syncthreads();                               // sW complete (e == 0) / sxj staged (e > 0)
        const float* wbase = sm.sW + rg16 * kSJ + jq * 32;
        const float* xbase = sm.sxj + (jq * 32) * kXJ + dg * 8;
#pragma unroll 2
        for (int jj = 0; jj < 8; ++jj) {
          float wv[4][4];
#pragma unroll
          for (int a = 0; a < 4; ++a) {
            const float4 w4 = *reinterpret_cast<const float4*>(wbase + a * 16 * kSJ + jj * 4);
            wv[a][0] = w4.x; wv[a][1] = w4.y; wv[a][2] = w4.z; wv[a][3] = w4.w;
            if (e == 0) swp[a] += (w4.x + w4.y) + (w4.z + w4.w);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 x0 = *reinterpret_cast<const float4*>(xbase + (jj * 4 + q) * kXJ);
            const float4 x1 = *reinterpret_cast<const float4*>(xbase + (jj * 4 + q) * kXJ + 4);
            const float xv[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
              for (int d = 0; d < 8; ++d) gacc[e][a][d] = fmaf(wv[a][q], xv[d], gacc[e][a][d]);
          }
        }
      }
    }
  }
  // ---- block results
  if (WITH_GRAD) {
    const int N_ = N;
    for (int e = 0; e < NDP; ++e) {
      if (dpass0 + e * kGD >= Dx) break;
      __syncthreads();
      float* red = sm.sW;                               // [4 jq][64 rows][32 dims]
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        float* dst = red + ((jq * kTI) + rg16 + 16 * a) * kGD + dg * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(gacc[e][a][0], gacc[e][a][1], gacc[e][a][2], gacc[e][a][3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(gacc[e][a][4], gacc[e][a][5], gacc[e][a][6], gacc[e][a][7]);
        if (e == 0 && dg == 0) sm.sw[jq * kTI + rg16 + 16 * a] = swp[a];
      }
      __syncthreads();
      const int r = tid >> 2, d8 = (tid & 3) * 8, i = i0 + r;
      if (i < iend) {
        const float sumW = (sm.sw[r] + sm.sw[kTI + r]) + (sm.sw[2 * kTI + r] + sm.sw[3 * kTI + r]);
#pragma unroll
        for (int d = 0; d < 8; ++d) {
          const int dim = dpass0 + e * kGD + d8 + d;
          if (dim < Dx) {
            const float g = (red[r * kGD + d8 + d] + red[(kTI + r) * kGD + d8 + d]) +
                            (red[(2 * kTI + r) * kGD + d8 + d] + red[(3 * kTI + r) * kGD + d8 + d]);
            const size_t off = ((size_t)n * k + i) * Dx + dim;
            gpart[((size_t)js * N_) * k * Dx + off] = fmaf(__ldg(x + off) - __ldg(xg + dim), sumW, -g);
          }
        }
      }
    }
  }
  if (dpass0 == 0) {
    lsum = warp_sum(lsum);
    __syncthreads();
    if (tx == 0) sm.red[ty] = lsum;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kNT / 32; ++w) t += sm.red[w];
      loss_part[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
    }
  }
}

// grad[n][i][d] = sum_js gpart[js][n][i][d], rows [row0,row0+rows), dims [d0,d1)
__global__ void ndiv_grad_reduce_kernel(const float* __restrict__ gpart, int N, int k, int Dx, int JS, int row0, int rows,
                                        int d0, int d1, float* __restrict__ grad,
                                        const int* __restrict__ run_if) {
  if (run_if && *run_if == 0) return;
  const int nd = d1 - d0;
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)N * rows * nd) return;
  const int d = d0 + (int)(t % nd);
  const size_t ri = t / nd;
  const int i = row0 + (int)(ri % rows), n = (int)(ri / rows);
  const size_t off = ((size_t)n * k + i) * Dx + d;
  float a = 0.f;
  for (int js = 0; js < JS; ++js) a += gpart[(size_t)js * N * k * Dx + off];
  grad[off] = a;
}

__global__ void ndiv_finalize_kernel(const float* __restrict__ part, int n, float* __restrict__ loss,
                                     const int* __restrict__ run_if) {
  if (run_if && *run_if == 0) return;
  __shared__ float s[256];
  float t = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) t += part[i];
  s[threadIdx.x] = t;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = s[0];
}

// ---------------------------------------------------------------- small groups (k <= 32): one warp per group
// The reference's own call shape (train_gan.py:193-196: N = batch*(seq-1) groups of k = num_sample = 6 samples,
// Dx = 4, Dz = 2) and every GAN-training shape: a 64x128 pair tile would be >95 % padding.  Lane = row i; the group's
// rows sit in shared memory (odd row stride); row sums, loss and gradient in ONE launch, same arithmetic as the
// tiled exact path (direct differences, IEEE sqrt / divide, hinge in the reference's operation order).
constexpr int kSmallWarps = 4;
__global__ void __launch_bounds__(kSmallWarps * 32)
ndiv_small_kernel(const float* __restrict__ x, const float* __restrict__ z, int N, int k, int Dx, int Dz, float margin,
                  float* __restrict__ sx_out, float* __restrict__ sz_out, float* __restrict__ loss_part,
                  float* __restrict__ grad_x) {
  extern __shared__ float sm_small[];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * kSmallWarps + w;
  if (n >= N) return;                                   // warps are independent: only __syncwarp below
  const int ldx = Dx | 1, ldz = Dz | 1;
  float* xs = sm_small + (size_t)w * (32 * (ldx + ldz) + 32 * ldx + 64);
  float* zs = xs + 32 * ldx;
  float* gs = zs + 32 * ldz;                            // gradient accumulators, lane-private rows
  float* rx = gs + 32 * ldx;                            // 1 / sx_j
  float* rz = rx + 32;                                  // 1 / sz_j
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  for (int e = lane; e < k * Dx; e += 32) xs[(e / Dx) * ldx + e % Dx] = __ldg(xg + e);
  for (int e = lane; e < k * Dz; e += 32) zs[(e / Dz) * ldz + e % Dz] = __ldg(zg + e);
  __syncwarp();
  const bool live = lane < k;
  const float* xi = xs + lane * ldx;
  const float* zi = zs + lane * ldz;
  float sxi = 0.f, szi = 0.f;
  if (live) {
    for (int j = 0; j < k; ++j) {
      float ax = 0.f, az = 0.f;
      for (int d = 0; d < Dx; ++d) { const float t = xi[d] - xs[j * ldx + d]; ax = fmaf(t, t, ax); }
      for (int d = 0; d < Dz; ++d) { const float t = zi[d] - zs[j * ldz + d]; az = fmaf(t, t, az); }
      sxi += sqrtf(ax);
      szi += sqrtf(az);
    }
    sx_out[(size_t)n * k + lane] = sxi;
    sz_out[(size_t)n * k + lane] = szi;
    rx[lane] = 1.f / sxi;
    rz[lane] = 1.f / szi;
    for (int d = 0; d < Dx; ++d) gs[lane * ldx + d] = 0.f;
  }
  __syncwarp();
  float lsum = 0.f;
  if (live) {
    const float rxi = rx[lane], rzi = rz[lane];
    for (int j = 0; j < k; ++j) {
      float ax = 0.f, az = 0.f;
      for (int d = 0; d < Dx; ++d) { const float t = xi[d] - xs[j * ldx + d]; ax = fmaf(t, t, ax); }
      for (int d = 0; d < Dz; ++d) { const float t = zi[d] - zs[j * ldz + d]; az = fmaf(t, t, az); }
      const float ddx = sqrtf(ax), ddz = sqrtf(az);
      const float hij = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, rzi)), __fmul_rn(ddx, rxi));
      const float hji = __fsub_rn(__fmul_rn(margin, __fmul_rn(ddz, rz[j])), __fmul_rn(ddx, rx[j]));
      lsum += !(hij <= 0.f) ? hij : 0.f;                // NaN propagates (k == 1 -> 0/0, like the reference)
      if (grad_x) {
        const float wij = (hij > 0.f ? rxi : 0.f) + (hji > 0.f ? rx[j] : 0.f);
        const float wh = ddx > 0.f ? -wij / ddx : 0.f;  // torch.norm subgradient: 0 at d == 0
        for (int d = 0; d < Dx; ++d) gs[lane * ldx + d] = fmaf(wh, xi[d] - xs[j * ldx + d], gs[lane * ldx + d]);
      }
    }
  }
  lsum = warp_sum(lsum);
  if (lane == 0) loss_part[n] = lsum;
  __syncwarp();
  if (grad_x)
    for (int e = lane; e < k * Dx; e += 32) grad_x[(size_t)n * k * Dx + e] = gs[(e / Dx) * ldx + e % Dx];
}
static bool small_eligible(int k, int Dx, int Dz) { return k >= 1 && k <= 32 && Dx + Dz <= 64; }
static size_t small_smem(int Dx, int Dz) {
  return (size_t)kSmallWarps * (32 * ((Dx | 1) + (Dz | 1)) + 32 * (Dx | 1) + 64) * sizeof(float);
}

// plain row sums of one distance matrix (normaliser of compute_pair_distance's backward): one warp per row
__global__ void rowsum_simple_kernel(const float* __restrict__ v, int k, int C, float* __restrict__ rs) {
  const int n = blockIdx.y;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= k) return;
  const float* vg = v + (size_t)n * k * C;
  float s = 0.f;
  for (int j = lane; j < k; j += 32) {
    float acc = 0.f;
    for (int c = 0; c < C; ++c) {
      const float df = __ldg(vg + (size_t)i * C + c) - __ldg(vg + (size_t)j * C + c);
      acc = fmaf(df, df, acc);
    }
    s += sqrtf(acc);
  }
  s = warp_sum(s);
  if (lane == 0) rs[(size_t)n * k + i] = s;
}

// ---------------------------------------------------------------- public pairwise API
// compute_pairwise / compute_pair_distance (diversity.py:8-19): [N,k,C] -> [N,k,k].
// One warp per output row i; lanes stride j. HBM-bound on the k*k write.
__global__ void pairwise_kernel(const float* __restrict__ v, int k, int C, int normalise,
                                float* __restrict__ out) {
  const int n = blockIdx.y;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= k) return;
  const float* vg = v + (size_t)n * k * C;
  float* og = out + ((size_t)n * k + i) * k;
  float rs = 0.f;
  for (int j = lane; j < k; j += 32) {
    float acc = 0.f;
    for (int c = 0; c < C; ++c) {
      const float df = __ldg(vg + (size_t)i * C + c) - __ldg(vg + (size_t)j * C + c);
      acc = fmaf(df, df, acc);
    }
    const float d = sqrtf(acc);
    og[j] = d;
    rs += d;
  }
  if (normalise) {
    rs = warp_sum(rs);
    __syncwarp();
    for (int j = lane; j < k; j += 32) og[j] = og[j] / rs;
  }
}

// grad_v_i = sum_j (g'_ij + g'_ji) (v_i - v_j)/d_ij, g' = g (plain) or g / s_row (normalised,
// normaliser detached).  One warp per row i.
__global__ void pairwise_bwd_kernel(const float* __restrict__ v, const float* __restrict__ go, int k, int C,
                                    int normalise, const float* __restrict__ rowsum, float* __restrict__ gv) {
  const int n = blockIdx.y;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* vg = v + (size_t)n * k * C;
  const float* gg = go + (size_t)n * k * k;
  const float* s_rowsum = rowsum + (size_t)n * k;
  const int i = blockIdx.x * warps + w;
  if (i >= k) return;
  for (int c0 = 0; c0 < C; c0 += 32) {
    const int c = c0 + lane;
    float g = 0.f;
    for (int j = 0; j < k; ++j) {
      float acc = 0.f;
      for (int cc = lane; cc < C; cc += 32) {
        const float df = __ldg(vg + (size_t)i * C + cc) - __ldg(vg + (size_t)j * C + cc);
        acc = fmaf(df, df, acc);
      }
      acc = warp_sum(acc);
      const float d = sqrtf(acc);
      float gij = __ldg(gg + (size_t)i * k + j), gji = __ldg(gg + (size_t)j * k + i);
      if (normalise) { gij /= s_rowsum[i]; gji /= s_rowsum[j]; }
      const float coef = d > 0.f ? (gij + gji) / d : 0.f;
      if (c < C) g = fmaf(coef, __ldg(vg + (size_t)i * C + c) - __ldg(vg + (size_t)j * C + c), g);
    }
    if (c < C) gv[((size_t)n * k + i) * C + c] = g;
  }
}

}  // namespace ndp

using namespace ndp;

namespace ndp {
// ---- tensor-core Gram fast path (k2_ndiv_tc.cu)
bool gram_eligible(int N, int k, int Dx, int Dz, int row0);
size_t gram_ws_bytes(int N, int k, int Dx, int Dz);
int gram_loss_parts(int N, int k);
int gram_js(int N, int k);
int gram_prepare(const float* x, const float* z, int N, int k, int Dx, int Dz, void* ws, cudaStream_t s);
int gram_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows, void* ws, float* sx,
                 float* sz, float* psx, float* psz, cudaStream_t s);
int gram_loss_grad(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx, int Dz,
                   float margin, int row0, int rows, void* ws, float* loss_part, float* gpart, float* loss, float* grad_x,
                   cudaStream_t s);

// j-split plan: depends on (N, k, Dx) only so that row-blocked (sharded) calls reproduce the full call bit for bit
struct NdivPlan { int ntiles, tps, JS; };
static NdivPlan ndiv_plan(int N, int k, int Dx) {
  NdivPlan p;
  p.ntiles = (k + kTJ - 1) / kTJ;
  if (p.ntiles < 1) p.ntiles = 1;
  const long long rowblocks = (long long)(N > 0 ? N : 1) * ((k + kTI - 1) / kTI > 0 ? (k + kTI - 1) / kTI : 1);
  long long want = 592 / rowblocks;                    // ~4 blocks per SM
  if (want < 1) want = 1;
  if (want > p.ntiles) want = p.ntiles;
  // keep the gradient partials below 64 MB
  const long long per = (long long)(N > 0 ? N : 1) * (k > 0 ? k : 1) * (Dx > 0 ? Dx : 1) * 4;
  while (want > 1 && want * per > (64ll << 20)) --want;
  p.tps = (int)((p.ntiles + want - 1) / want);
  p.JS = (p.ntiles + p.tps - 1) / p.tps;
  return p;
}
// workspace: [loss partials | partial sums (row sums or gradients) of the column splits | Gram-path images]
struct NdivWs { size_t part, psum, gram, total; };
static NdivWs ndiv_ws(int N, int k, int Dx, int Dz) {
  const NdivPlan p = ndiv_plan(N, k, Dx);
  const bool tc = gram_eligible(N, k, Dx, Dz, 0);
  const int JS = tc && gram_js(N, k) > p.JS ? gram_js(N, k) : p.JS;
  NdivWs w;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t rb = (size_t)((k + kTI - 1) / kTI + 1);
  size_t parts = (size_t)(N > 0 ? N : 1) * rb * p.JS;
  if (tc && (size_t)gram_loss_parts(N, k) > parts) parts = (size_t)gram_loss_parts(N, k);
  w.part = al(parts * sizeof(float));
  const size_t ps = JS > 1 ? (size_t)2 * JS * N * k * sizeof(float) : 0;
  const size_t gp = JS > 1 ? (size_t)JS * N * k * Dx * sizeof(float) : 0;
  w.psum = al(ps > gp ? ps : gp);
  w.gram = tc ? gram_ws_bytes(N, k, Dx, Dz) : 0;
  w.total = w.part + w.psum + w.gram + 256;
  return w;
}
}  // namespace ndp

extern "C" int ndp_ndiv_workspace_bytes(int N, int k, int Dx, int Dz, size_t* bytes) {
  NDP_CHECK_ARG(bytes && N >= 0 && k >= 0 && Dx >= 1 && Dz >= 1, "ndiv: bad N/k/Dx/Dz");
  *bytes = ndiv_ws(N, k, Dx, Dz).total;
  return NDP_OK;
}

static int ndiv_args_ok(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows) {
  NDP_CHECK_ARG(N >= 0 && k >= 0 && Dx >= 1 && Dz >= 1, "ndiv: bad shape N=%d k=%d Dx=%d Dz=%d", N, k, Dx, Dz);
  NDP_CHECK_ARG(N <= 65535, "ndiv: N must be <= 65535 per call");
  NDP_CHECK_ARG(row0 >= 0 && rows >= 0 && row0 + rows <= k, "ndiv: bad row block [%d,+%d) of %d", row0, rows, k);
  NDP_CHECK_ARG(N == 0 || k == 0 || (x && z), "ndiv: null x/z");
  return NDP_OK;
}

// exact direct-difference pass 1; run_if (nullable, device) gates the whole pass
static int exact_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows, float* sx,
                         float* sz, char* ws, const NdivWs& w, const int* run_if, cudaStream_t s) {
  const NdivPlan p = ndiv_plan(N, k, Dx);
  dim3 grid((rows + kTI - 1) / kTI, p.JS, N);
  const size_t smem1 = (size_t)(kSD * kSI + kSD * kSJ + 4 * kTI) * sizeof(float);
  NDP_CUDA(cudaFuncSetAttribute(ndiv_rowsum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
  if (p.JS == 1) {
    ndiv_rowsum_kernel<<<grid, kNT, smem1, s>>>(x, z, k, Dx, Dz, row0, rows, p.tps, sx, sz, run_if);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  float* psx = reinterpret_cast<float*>(ws + w.part);
  float* psz = psx + (size_t)p.JS * N * k;
  ndiv_rowsum_kernel<<<grid, kNT, smem1, s>>>(x, z, k, Dx, Dz, row0, rows, p.tps, psx, psz, run_if);
  NDP_LAUNCH_CHECK();
  ndiv_rowsum_reduce_kernel<<<(N * rows + 255) / 256, 256, 0, s>>>(psx, psz, N, k, p.JS, row0, rows, sx, sz, run_if);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// tensor-core pass 1 on prepared images (gram_prepare): ONE launch, final row sums (the column splits are combined by
// the last CTA of each row tile)
static int tc_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows, float* sx,
                      float* sz, char* b, const NdivWs& w, cudaStream_t s) {
  void* gws = b + w.part + w.psum;
  const int JS = gram_js(N, k);
  float* psx = reinterpret_cast<float*>(b + w.part);
  float* psz = psx + (size_t)JS * N * k;
  return gram_rowsums(x, z, N, k, Dx, Dz, row0, rows, gws, sx, sz, psx, psz, s);
}

// tensor-core pass 2 on prepared images: ONE launch, final loss and gradient
static int tc_loss_grad(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx, int Dz,
                        float margin, int row0, int rows, float* loss, float* grad_x, char* b, const NdivWs& w,
                        cudaStream_t s) {
  void* gws = b + w.part + w.psum;
  float* part = reinterpret_cast<float*>(b);
  float* gpart = reinterpret_cast<float*>(b + w.part);
  return gram_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, row0, rows, gws, part, gpart, loss, grad_x, s);
}

extern "C" int ndp_ndiv_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz,
                                int row0, int rows, float* sx, float* sz, void* ws, size_t ws_bytes, void* stream) {
  if (int rc = ndiv_args_ok(x, z, N, k, Dx, Dz, row0, rows)) return rc;
  if (int rc = device_check()) return rc;
  if (N == 0 || rows == 0) return NDP_OK;
  NDP_CHECK_ARG(sx && sz, "ndiv: null sx/sz");
  cudaStream_t s = (cudaStream_t)stream;
  const NdivWs w = ndiv_ws(N, k, Dx, Dz);
  const bool tc = gram_eligible(N, k, Dx, Dz, row0);
  const bool need_ws = tc || ndiv_plan(N, k, Dx).JS > 1;
  if (need_ws && (!ws || ws_bytes < w.total)) { set_error("ndiv: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  char* b = reinterpret_cast<char*>(ws);
  if (tc) {
    // Gram form on tensor cores; ill-conditioned pairs are recomputed exactly inside the kernel (k2_ndiv_tc.cu)
    void* gws = b + w.part + w.psum;
    if (int rc = gram_prepare(x, z, N, k, Dx, Dz, gws, s)) return rc;
    return tc_rowsums(x, z, N, k, Dx, Dz, row0, rows, sx, sz, b, w, s);
  }
  return exact_rowsums(x, z, N, k, Dx, Dz, row0, rows, sx, sz, b, w, nullptr, s);
}

template <int NDP, bool WG>
static int launch_loss_grad(dim3 grid, cudaStream_t s, const float* x, const float* z, const float* sx,
                            const float* sz, int k, int Dx, int Dz, float margin, int row0, int rows, int tps,
                            int dpass0, float* part, float* gpart, const int* run_if) {
  NDP_CUDA(cudaFuncSetAttribute(ndiv_loss_grad_kernel<NDP, WG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(NdivSmem2)));
  ndiv_loss_grad_kernel<NDP, WG><<<grid, kNT, sizeof(NdivSmem2), s>>>(x, z, sx, sz, k, Dx, Dz, margin, row0, rows, tps,
                                                                      dpass0, part, gpart, run_if);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

static int exact_loss_grad(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx, int Dz,
                           float margin, int row0, int rows, float* loss, float* grad_x, char* ws, const NdivWs& w,
                           const int* run_if, cudaStream_t s) {
  const NdivPlan p = ndiv_plan(N, k, Dx);
  dim3 grid((rows + kTI - 1) / kTI, p.JS, N);
  const size_t nparts = (size_t)grid.x * grid.y * grid.z;
  float* part = reinterpret_cast<float*>(ws);
  float* gpart = p.JS > 1 ? reinterpret_cast<float*>(ws + w.part) : grad_x;
  if (!grad_x) {
    if (int rc = launch_loss_grad<1, false>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, p.tps, 0, part, nullptr, run_if)) return rc;
  } else {
    const int per_launch = 4 * kGD;
    for (int dp = 0; dp < Dx; dp += per_launch) {
      const int passes = ((Dx - dp < per_launch ? Dx - dp : per_launch) + kGD - 1) / kGD;
      int rc;
      if (passes <= 1) rc = launch_loss_grad<1, true>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, p.tps, dp, part, gpart, run_if);
      else if (passes <= 2) rc = launch_loss_grad<2, true>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, p.tps, dp, part, gpart, run_if);
      else rc = launch_loss_grad<4, true>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, p.tps, dp, part, gpart, run_if);
      if (rc) return rc;
    }
    if (p.JS > 1) {
      const size_t tot = (size_t)N * rows * Dx;
      ndiv_grad_reduce_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(gpart, N, k, Dx, p.JS, row0, rows, 0, Dx, grad_x, run_if);
      NDP_LAUNCH_CHECK();
    }
  }
  ndiv_finalize_kernel<<<1, 256, 0, s>>>(part, (int)nparts, loss, run_if);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_ndiv_loss_grad(const float* x, const float* z, const float* sx, const float* sz,
                                  int N, int k, int Dx, int Dz, float margin, int row0, int rows,
                                  float* loss, float* grad_x, void* ws, size_t ws_bytes, void* stream) {
  if (int rc = ndiv_args_ok(x, z, N, k, Dx, Dz, row0, rows)) return rc;
  if (int rc = device_check()) return rc;
  NDP_CHECK_ARG(loss, "ndiv: null loss");
  cudaStream_t s = (cudaStream_t)stream;
  if (N == 0 || rows == 0) {
    NDP_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), s));
    return NDP_OK;
  }
  NDP_CHECK_ARG(sx && sz, "ndiv: null sx/sz");
  const NdivWs w = ndiv_ws(N, k, Dx, Dz);
  if (!ws || ws_bytes < w.total) { set_error("ndiv: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  char* b = reinterpret_cast<char*>(ws);
  if (gram_eligible(N, k, Dx, Dz, row0)) {
    void* gws = b + w.part + w.psum;
    if (int rc = gram_prepare(x, z, N, k, Dx, Dz, gws, s)) return rc;
    return tc_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, row0, rows, loss, grad_x, b, w, s);
  }
  return exact_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, row0, rows, loss, grad_x, b, w, nullptr, s);
}

// Both passes over all rows in one call (the single-GPU autograd step): operands are pre-packed once.
extern "C" int ndp_ndiv_forward_backward(const float* x, const float* z, int N, int k, int Dx, int Dz, float margin,
                                         float* sx, float* sz, float* loss, float* grad_x, void* ws, size_t ws_bytes,
                                         void* stream) {
  if (int rc = ndiv_args_ok(x, z, N, k, Dx, Dz, 0, k)) return rc;
  if (int rc = device_check()) return rc;
  NDP_CHECK_ARG(loss, "ndiv: null loss");
  cudaStream_t s = (cudaStream_t)stream;
  if (N == 0 || k == 0) {
    NDP_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), s));
    return NDP_OK;
  }
  NDP_CHECK_ARG(sx && sz, "ndiv: null sx/sz");
  const NdivWs w = ndiv_ws(N, k, Dx, Dz);
  if (!ws || ws_bytes < w.total) { set_error("ndiv: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  char* b = reinterpret_cast<char*>(ws);
  if (small_eligible(k, Dx, Dz)) {
    float* part = reinterpret_cast<float*>(b);          // N loss partials (ndiv_ws reserves >= 2N)
    static std::atomic<unsigned long long> attr_devs{0};
    if (first_use_on_device(attr_devs)) {
      NDP_CUDA(cudaFuncSetAttribute(ndiv_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)small_smem(62, 2)));
    }
    ndiv_small_kernel<<<(N + kSmallWarps - 1) / kSmallWarps, kSmallWarps * 32, small_smem(Dx, Dz), s>>>(
        x, z, N, k, Dx, Dz, margin, sx, sz, part, grad_x);
    NDP_LAUNCH_CHECK();
    ndiv_finalize_kernel<<<1, 256, 0, s>>>(part, N, loss, nullptr);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  if (gram_eligible(N, k, Dx, Dz, 0)) {               // three launches: pre-pack, row sums, loss + gradient
    void* gws = b + w.part + w.psum;
    if (int rc = gram_prepare(x, z, N, k, Dx, Dz, gws, s)) return rc;
    if (int rc = tc_rowsums(x, z, N, k, Dx, Dz, 0, k, sx, sz, b, w, s)) return rc;
    return tc_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, 0, k, loss, grad_x, b, w, s);
  }
  if (int rc = exact_rowsums(x, z, N, k, Dx, Dz, 0, k, sx, sz, b, w, nullptr, s)) return rc;
  return exact_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, 0, k, loss, grad_x, b, w, nullptr, s);
}

// ---- forward / backward split of the operator (SURVEY.md 8b names): what a torch.autograd.Function binds --------
// forward: loss[0] and the saved normalisers sx, sz [N,k]
extern "C" int ndp_ndiv_forward(const float* x, const float* z, int N, int k, int Dx, int Dz, float margin, float* loss,
                                float* sx, float* sz, void* ws, size_t ws_bytes, void* stream) {
  return ndp_ndiv_forward_backward(x, z, N, k, Dx, Dz, margin, sx, sz, loss, nullptr, ws, ws_bytes, stream);
}
__global__ void ndiv_scale_kernel(float* __restrict__ g, const float* __restrict__ scale, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) g[i] *= scale[0];
}
// d loss / d dz_ij = margin * M_ij / sz_i  (M_ij = hinge active; normalisers detached, diversity.py:17-18): one warp per
// row i writes row i of G [N,k,k]; ndp_pairwise_backward's kernel then turns G into d loss / d z.  Secondary path
// (the reference call site never differentiates the latent codes): direct differences, k x k scratch.
__global__ void ndiv_hinge_dz_kernel(const float* __restrict__ x, const float* __restrict__ z, const float* __restrict__ sx,
                                     const float* __restrict__ sz, int k, int Dx, int Dz, float margin,
                                     float* __restrict__ G) {
  const int n = blockIdx.y;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= k) return;
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  const float sxi = __ldg(sx + (size_t)n * k + i), szi = __ldg(sz + (size_t)n * k + i);
  for (int j = lane; j < k; j += 32) {
    float ax = 0.f, az = 0.f;
    for (int c = 0; c < Dx; ++c) { const float df = __ldg(xg + (size_t)i * Dx + c) - __ldg(xg + (size_t)j * Dx + c); ax = fmaf(df, df, ax); }
    for (int c = 0; c < Dz; ++c) { const float df = __ldg(zg + (size_t)i * Dz + c) - __ldg(zg + (size_t)j * Dz + c); az = fmaf(df, df, az); }
    const float h = __fsub_rn(__fmul_rn(__fdiv_rn(sqrtf(az), szi), margin), __fdiv_rn(sqrtf(ax), sxi));   // diversity.py:40
    G[((size_t)n * k + i) * k + j] = h > 0.f ? __fdiv_rn(margin, szi) : 0.f;
  }
}

extern "C" int ndp_ndiv_gradz_workspace_bytes(int N, int k, int Dx, int Dz, size_t* bytes) {
  NDP_CHECK_ARG(bytes && N >= 0 && k >= 0 && Dx >= 1 && Dz >= 1, "ndiv: bad N/k/Dx/Dz");
  *bytes = (ndiv_ws(N, k, Dx, Dz).total + 255) / 256 * 256 + (size_t)N * k * k * sizeof(float);
  return NDP_OK;
}

// backward: grad_x = grad_loss[0] * d loss / d x from the saved normalisers (grad_loss: device scalar, NULL = 1).
// grad_z (nullable) = grad_loss[0] * d loss / d z — what autograd gives a caller that sets codes.requires_grad
// (diversity.py:36-41); needs ws of ndp_ndiv_gradz_workspace_bytes() (an [N,k,k] scratch behind the usual workspace).
// loss_scratch: device float the kernels may overwrite (the loss is recomputed on the way).
extern "C" int ndp_ndiv_backward(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx,
                                 int Dz, float margin, const float* grad_loss, float* grad_x, float* grad_z,
                                 float* loss_scratch, void* ws, size_t ws_bytes, void* stream) {
  NDP_CHECK_ARG((grad_x || grad_z) && loss_scratch, "ndp_ndiv_backward: null grad_x and grad_z / loss_scratch");
  cudaStream_t s = (cudaStream_t)stream;
  if (grad_x) {
    if (int rc = ndp_ndiv_loss_grad(x, z, sx, sz, N, k, Dx, Dz, margin, 0, k, loss_scratch, grad_x, ws, ws_bytes, stream))
      return rc;
    if (grad_loss && N > 0 && k > 0) {
      const size_t n = (size_t)N * k * Dx;
      ndiv_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(grad_x, grad_loss, n);
      NDP_LAUNCH_CHECK();
    }
  }
  if (grad_z && N > 0 && k > 0) {
    if (int rc = ndiv_args_ok(x, z, N, k, Dx, Dz, 0, k)) return rc;
    if (int rc = device_check()) return rc;
    NDP_CHECK_ARG(sx && sz, "ndiv: null sx/sz");
    size_t need = 0;
    ndp_ndiv_gradz_workspace_bytes(N, k, Dx, Dz, &need);
    if (!ws || ws_bytes < need) { set_error("ndiv backward (grad_z): workspace %zu < %zu", ws_bytes, need); return NDP_ERR_WORKSPACE; }
    float* G = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + (need - (size_t)N * k * k * sizeof(float)));
    dim3 grid((k + 7) / 8, N);
    ndiv_hinge_dz_kernel<<<grid, 256, 0, s>>>(x, z, sx, sz, k, Dx, Dz, margin, G);
    NDP_LAUNCH_CHECK();
    pairwise_bwd_kernel<<<grid, 256, 0, s>>>(z, G, k, Dz, 0, z, grad_z);
    NDP_LAUNCH_CHECK();
    if (grad_loss) {
      const size_t n = (size_t)N * k * Dz;
      ndiv_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(grad_z, grad_loss, n);
      NDP_LAUNCH_CHECK();
    }
  }
  return NDP_OK;
}

extern "C" int ndp_pairwise(const float* v, int N, int k, int C, int normalise, float* out, void* stream) {
  NDP_CHECK_ARG(N >= 0 && k >= 0 && C >= 1 && N <= 65535, "pairwise: bad shape");
  if (int rc = device_check()) return rc;
  if (N == 0 || k == 0) return NDP_OK;
  NDP_CHECK_ARG(v && out, "pairwise: null pointer");
  dim3 grid((k + 7) / 8, N);
  pairwise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(v, k, C, normalise, out);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_pairwise_backward(const float* v, const float* grad_out, int N, int k, int C,
                                     int normalise, float* rowsum_ws, float* grad_v, void* stream) {
  NDP_CHECK_ARG(N >= 0 && k >= 0 && C >= 1 && N <= 65535, "pairwise_backward: bad shape");
  NDP_CHECK_ARG(!normalise || rowsum_ws, "pairwise_backward: normalise needs rowsum_ws [N,k]");
  if (int rc = device_check()) return rc;
  if (N == 0 || k == 0) return NDP_OK;
  NDP_CHECK_ARG(v && grad_out && grad_v, "pairwise_backward: null pointer");
  dim3 grid((k + 7) / 8, N);
  if (normalise) {
    rowsum_simple_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(v, k, C, rowsum_ws);
    NDP_LAUNCH_CHECK();
  }
  pairwise_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(v, grad_out, k, C, normalise,
                                                             normalise ? rowsum_ws : v, grad_v);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
