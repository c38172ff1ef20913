// K4 — the GAN training step around the ndiv loss (train_gan.py:159-203), device resident:
//
//   a_hat  = G([codes | noise])                                    train_gan.py:159-165   (gan.py:79-86)
//   repeat d_steps:  D_loss = BCE(D(a, codes), 1) + BCE(D(a_hat, codes), 0);  Adam(D)      :172-184
//   G_loss = BCE(D(a_hat, codes), 1)                               :187-190  (discriminator AFTER its update)
//   div    = compute_pairwise_divergence(a_hat[F,S,4], noise[F,S,nd])   :193-196  (K2, k2_ndiv.cu)
//   total  = G_loss + div_factor * div;  Adam(G)                   :198-203
//
// Everything is stream-ordered with no host synchronisation, so a caller may capture one step in a CUDA graph.
// This file is the exact-arithmetic (fp32 FFMA) implementation: every dense layer, its input gradient and its
// weight gradient go through ONE smem-tiled kernel template (gemm_kernel<MODE,TJ>) whose operand loaders
// express the three products without materialising a concatenation, a repeat_interleave or a transpose:
//
//   FWD  Y[m,n]   = act( sum_k X(m,k) W[n,k] + b[n] )               X(m,k) = [P0[m/div0] | P1[m/div1]]
//   DX   dP'[m,k] = ( sum_n dP[m,n] W[n,k0+k] ) * act'(H[m,k]) (+ s * addend[m,k])
//   DW   dW[n,k]  = sum_m dP[m,n] X(m,k),  db[n] = sum_m dP[m,n]    rows split over blockIdx.z, partials summed
//                                                                   in a fixed order by the Adam kernel
//
// Algebra the reference leaves on the table and this step uses: the "real" discriminator batch is S identical
// copies of F distinct rows (repeat_interleave at :139-141), and BCE is a mean, so loss and gradients are
// evaluated on the F distinct rows only; torch.cat / repeat_interleave inputs are read through (row / div)
// index maps.  Weight gradients are deterministic (no atomics).
#include <math.h>
#include <stdlib.h>
#include <vector>
#include "common.cuh"
#include "tc_common.cuh"

namespace ndp {
namespace gan {

enum { FWD = 0, DX = 1, DW = 2 };
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_LEAKY = 2 };
constexpr int kTI = 64, kTC = 16, kThreads = 256;
constexpr float kLeaky = 0.01f;   // F.leaky_relu default negative_slope (gan.py:103-105)

struct RowSrc {   // logical [M x (k0+k1)] matrix assembled from two row-indexed sources
  const float* p0; int k0, ld0, div0; unsigned mul0;
  const float* p1; int k1, ld1, div1; unsigned mul1;
  // r / div for r < 2^24, div < 2^8 via one multiply-high (mul = floor(2^32/div) + 1); div == 1 passes through
  static __device__ __forceinline__ int fdiv(int r, int div, unsigned mul) {
    return div == 1 ? r : (int)__umulhi((unsigned)r, mul);
  }
  __device__ __forceinline__ const float* ptr(int r, int k) const {
    return k < k0 ? p0 + (size_t)fdiv(r, div0, mul0) * ld0 + k : p1 + (size_t)fdiv(r, div1, mul1) * ld1 + (k - k0);
  }
};
static RowSrc row_src(const float* p0, int k0, int div0, const float* p1 = nullptr, int k1 = 0, int div1 = 1) {
  RowSrc r; r.p0 = p0; r.k0 = k0; r.ld0 = k0; r.div0 = div0; r.p1 = p1; r.k1 = k1; r.ld1 = k1; r.div1 = div1;
  r.mul0 = (unsigned)((1ull << 32) / (unsigned)div0 + 1); r.mul1 = (unsigned)((1ull << 32) / (unsigned)div1 + 1);
  return r;
}

struct GemmP {
  int I, J, Cn;            // output rows, output cols, contraction length
  RowSrc x;                // FWD: A operand; DW: B operand
  const float* W; int ldw, wk0;
  const float* b; int act;
  const float* dP; int ldp;
  const float* hprev; int ldh, act_prev;
  const float* addend; float add_scale;
  float* out; int ldo;
  int rows_per_split; size_t split_stride; float* db_out;
  const unsigned* amax_in;   // tensor path: max |dP| of the A operand (bits of a non-negative float), NULL = unscaled
  unsigned* amax_out;        // DX: max |output| is accumulated here (atomicMax on the float bits), may be NULL
  int use_tc;
  const uint8_t* b_img;      // tensor path, FWD / DX: pre-split weight image (pack_weights_kernel), NULL = split on the fly
  int b_chunks;              // contraction chunks per 128-row tile of that image
};

__device__ __forceinline__ float act_fwd(float v, int act) {
  if (act == ACT_RELU) return fmaxf(v, 0.f);
  if (act == ACT_LEAKY) return v > 0.f ? v : v * kLeaky;
  return v;
}
// derivative from the saved post-activation value (sign(h) == sign(pre-activation) for both)
__device__ __forceinline__ float act_bwd(float h, int act) {
  if (act == ACT_RELU) return h > 0.f ? 1.f : 0.f;
  if (act == ACT_LEAKY) return h > 0.f ? 1.f : kLeaky;
  return 1.f;
}

// 4-byte async copy global -> shared; src_bytes == 0 zero-fills (out-of-range elements)
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int kStages = 4;
static int num_sms_cached() {
  static int n = 0;
  if (!n) n = num_sms();
  return n;
}

template <int MODE, int TI, int TJ>
__global__ void __launch_bounds__(kThreads) gemm_kernel(const GemmP p) {
  constexpr int RI = TI / 16, RJ = TJ / 16;
  __shared__ __align__(16) float As[kStages][kTC][TI + 4];
  __shared__ __align__(16) float Bs[kStages][kTC][TJ + 4];
  const int t = threadIdx.x;
  const int i0 = blockIdx.x * TI, j0 = blockIdx.y * TJ;
  int c_begin = 0, c_end = p.Cn;
  if (MODE == DW) {
    c_begin = blockIdx.z * p.rows_per_split;
    c_end = min(p.Cn, c_begin + p.rows_per_split);
  }
  const int ti = t >> 4, tj = t & 15;
  float acc[RI][RJ];
#pragma unroll
  for (int a = 0; a < RI; a++)
#pragma unroll
    for (int b = 0; b < RJ; b++) acc[a][b] = 0.f;
  float bsum = 0.f;

  // the chunk [c0, c0+kTC) of both operands goes global -> shared as element-wise async copies (the layouts are
  // transposed / gathered on the way), kStages-1 chunks in flight
  auto issue = [&](int stage, int c0) {
#pragma unroll
    for (int q = 0; q < TI * kTC / kThreads; q++) {
      const int e = t + q * kThreads;
      int i, c;
      if (MODE == DW) { i = e % TI; c = e / TI; } else { c = e & (kTC - 1); i = e >> 4; }
      const int gi = i0 + i, gc = c0 + c;
      const bool ok = gi < p.I && gc < c_end;
      const float* src = MODE == FWD ? p.x.p0 : p.dP;
      if (ok) {
        if (MODE == FWD) src = p.x.ptr(gi, gc);
        else if (MODE == DX) src = p.dP + (size_t)gi * p.ldp + gc;
        else src = p.dP + (size_t)gc * p.ldp + gi;
      }
      cp_async4(&As[stage][c][i], src, ok);
    }
    for (int e = t; e < TJ * kTC; e += kThreads) {
      int j, c;
      if (MODE == FWD) { c = e & (kTC - 1); j = e >> 4; } else { j = e % TJ; c = e / TJ; }
      const int gj = j0 + j, gc = c0 + c;
      const bool ok = gj < p.J && gc < c_end;
      const float* src = MODE == DW ? p.x.p0 : p.W;
      if (ok) {
        if (MODE == FWD) src = p.W + (size_t)gj * p.ldw + gc;
        else if (MODE == DX) src = p.W + (size_t)gc * p.ldw + p.wk0 + gj;
        else src = p.x.ptr(gc, gj);
      }
      cp_async4(&Bs[stage][c][j], src, ok);
    }
  };

  const int nchunks = (c_end - c_begin + kTC - 1) / kTC;
#pragma unroll
  for (int st = 0; st < kStages - 1; st++) {
    if (st < nchunks) issue(st, c_begin + st * kTC);
    cp_async_commit();
  }
  for (int n = 0; n < nchunks; n++) {
    const int st = n % kStages;
    cp_async_wait<kStages - 2>();
    __syncthreads();          // chunk n landed for every thread; everyone is done reading stage (n-1) % kStages
    if (n + kStages - 1 < nchunks) issue((n + kStages - 1) % kStages, c_begin + (n + kStages - 1) * kTC);
    cp_async_commit();
    if (MODE == DW && blockIdx.y == 0 && t < TI) {
#pragma unroll
      for (int c = 0; c < kTC; c++) bsum += As[st][c][t];
    }
#pragma unroll
    for (int c = 0; c < kTC; c++) {
      float a[RI];
      if (RI == 4) {
        const float4 av = *reinterpret_cast<const float4*>(&As[st][c][ti * 4]);
        a[0] = av.x; a[1 % RI] = av.y; a[2 % RI] = av.z; a[3 % RI] = av.w;
      } else {
        a[0] = As[st][c][ti];
      }
      float b[RJ];
      if (RJ == 4) {
        const float4 bv = *reinterpret_cast<const float4*>(&Bs[st][c][tj * 4]);
        b[0] = bv.x; b[1 % RJ] = bv.y; b[2 % RJ] = bv.z; b[3 % RJ] = bv.w;
      } else {
        b[0] = Bs[st][c][tj];
      }
#pragma unroll
      for (int x = 0; x < RI; x++)
#pragma unroll
        for (int y = 0; y < RJ; y++) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
    }
  }

  float* out = p.out;
  if (MODE == DW) out += (size_t)blockIdx.z * p.split_stride;
  float vmax = 0.f;
#pragma unroll
  for (int x = 0; x < RI; x++) {
    const int gi = i0 + ti * RI + x;
    if (gi >= p.I) continue;
#pragma unroll
    for (int y = 0; y < RJ; y++) {
      const int gj = j0 + tj * RJ + y;
      if (gj >= p.J) continue;
      float v = acc[x][y];
      if (MODE == FWD) {
        v = act_fwd(v + __ldg(p.b + gj), p.act);
      } else if (MODE == DX) {
        if (p.hprev) v *= act_bwd(__ldg(p.hprev + (size_t)gi * p.ldh + gj), p.act_prev);
        if (p.addend) v = fmaf(p.add_scale, __ldg(p.addend + (size_t)gi * p.J + gj), v);
      }
      out[(size_t)gi * p.ldo + gj] = v;
      if (MODE == DX) vmax = fmaxf(vmax, fabsf(v));
    }
  }
  if (MODE == DX && p.amax_out) {
    for (int o = 16; o; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if ((t & 31) == 0 && vmax > 0.f) atomicMax(p.amax_out, __float_as_uint(vmax));
  }
  if (MODE == DW && blockIdx.y == 0 && t < TI && i0 + t < p.I)
    p.db_out[(size_t)blockIdx.z * p.split_stride + i0 + t] = bsum;
}

// ---- tensor-core variant of the three products (tcgen05, kind::f16, split-fp16 operands) ---------------------
// Same operand semantics as gemm_kernel, 128 x 128 output tile per block, contraction in chunks of 64.  The block's
// threads read the fp32 operands (through the same index maps), split every value into fp16 hi + fp16 lo (exact
// residual) and write both halves straight into the K-major core-matrix image the MMA reads (the layout of
// k2_ndiv_tc.cu: slab of K=16 = 4096 B, 8x16 B core matrices, LBO 2048, SBO 128); one thread issues
// hi*hi + lo*hi + hi*lo per slab (22-bit products, fp32 accumulation in TMEM) while the block stages the next chunk
// into the other buffer.  Gradient operands (dP ~ 1/batch) are multiplied by a power of two taken from the running
// max |dP| of the producing kernel so that their lo halves stay normal fp16 numbers; the epilogue divides it out.
using namespace ndp::tc;
constexpr int kTcK = 32;                 // contraction elements per chunk (2 slabs of K = 16)
constexpr int kTcU = kTcK / 16;          // (row, 8-group) units per thread and operand: 128 * kTcK / 8 / 256
constexpr int kTcHalf = 128 * kTcK * 2;  // one 128 x kTcK fp16 image
constexpr int kTcOp = 2 * kTcHalf;       // hi + lo
constexpr int kTcStage = 2 * kTcOp;      // A + B
constexpr int kTcTile = 128 * 132 * 4;   // fp32 epilogue tile (aliases the operand buffers)
constexpr int kTcBuf = 2 * kTcStage > kTcTile ? 2 * kTcStage : kTcTile;
constexpr int kTcSmem = kTcBuf + 1024 + 64;   // ~69 KB: two to three blocks per SM hide each other's load latency

constexpr float kUnitScale = 256.f;
__device__ __forceinline__ float pow2_scale(const unsigned* amax) {
  if (!amax) return kUnitScale;
  const float m = __uint_as_float(*amax);
  if (!(m > 0.f) || !(m < 3e38f)) return kUnitScale;
  int e;
  frexpf(m, &e);                          // m = f * 2^e, f in [0.5, 1)
  int sh = 9 - e;                         // m * 2^sh in [256, 512)
  sh = sh > 100 ? 100 : (sh < -100 ? -100 : sh);
  return ldexpf(1.f, sh);
}

// 8 consecutive contraction elements starting at `base` (nvalid of them in range), vectorised when aligned
__device__ __forceinline__ void load8(const float* base, int nvalid, float (&v)[8]) {
  if (nvalid >= 8 && ((reinterpret_cast<uintptr_t>(base) & 15) == 0)) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(base)), b = __ldg(reinterpret_cast<const float4*>(base) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = e < nvalid ? __ldg(base + e) : 0.f;
  }
}
// row r of a RowSrc, columns [c, c+8) clipped to c_end
__device__ __forceinline__ void load8_rowsrc(const RowSrc& x, int r, int c, int c_end, float (&v)[8]) {
  const int n = c_end - c;
  if (c + 8 <= x.k0) {
    load8(x.p0 + (size_t)RowSrc::fdiv(r, x.div0, x.mul0) * x.ld0 + c, n, v);
  } else if (c >= x.k0) {
    const int rem = x.k0 + x.k1 - c;
    load8(x.p1 + (size_t)RowSrc::fdiv(r, x.div1, x.mul1) * x.ld1 + (c - x.k0), n < rem ? n : rem, v);
  } else {
#pragma unroll
    for (int e = 0; e < 8; e++) v[e] = (e < n && c + e < x.k0 + x.k1) ? __ldg(x.ptr(r, c + e)) : 0.f;
  }
}

// (row, 8-group) unit of a 128 x 64 operand chunk -> hi / lo core-matrix rows
__device__ __forceinline__ void store_unit(uint8_t* op, int r, int g, const float (&v)[8], float scale) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const float a = v[2 * q] * scale, b = v[2 * q + 1] * scale;
    h[q] = pack_hi(a, b);
    l[q] = pack_lo(a, b, h[q]);
  }
  const uint32_t off = (uint32_t)((g >> 1) * 4096 + (g & 1) * 2048 + (r >> 3) * 128 + (r & 7) * 16);
  *reinterpret_cast<uint4*>(op + off) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(op + kTcHalf + off) = make_uint4(l[0], l[1], l[2], l[3]);
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 3) gemm_tc_kernel(const GemmP p) {
  extern __shared__ __align__(128) uint8_t tc_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int i0 = blockIdx.x * 128, j0 = blockIdx.y * 128;
  int c_begin = 0, c_end = p.Cn;
  if (MODE == DW) {
    c_begin = blockIdx.z * p.rows_per_split;
    c_end = min(p.Cn, c_begin + p.rows_per_split);
  }
  const int nchunks = (c_end - c_begin + kTcK - 1) / kTcK;
  float* sbsum = reinterpret_cast<float*>(tc_smem + kTcBuf);                 // [256]
  uint8_t* sbar = tc_smem + kTcBuf + 1024;
  const uint32_t bar0 = smem_u32(sbar);                                        // bar_done[2]
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(sbar + 32);
  const uint32_t barf = bar0 + 16;                                             // bar_full[2]: weight image landed
  if (tid == 0) {
    mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); mbar_init(barf, 1); mbar_init(barf + 8, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc128(smem_u32((const void*)tmem_slot));
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  // operands without a tracked maximum (inputs, activations, weights: O(1e-3 .. 10)) are multiplied by 2^8 so that
  // their lo halves are normal fp16 numbers (|v| must stay below 255, as in K1's tensor modes)
  const float a_scale = MODE == FWD ? kUnitScale : pow2_scale(p.amax_in);
  const float b_scale = kUnitScale;
  const int nb = (p.J - j0) <= 64 ? 64 : 128;
  const uint32_t idesc = idesc_f16(nb);

  // unit maps.  C: source contiguous along the contraction -> quarter-warps own 8 consecutive rows of one 8-group
  // (conflict-free 16 B stores).  R: source contiguous along the operand rows -> a warp owns 32 consecutive rows.
  constexpr bool A_C = MODE != DW, B_C = MODE == FWD;
  // FWD / DX: the B operand is a weight matrix, identical for every block -> split once per optimiser step by
  // pack_weights_kernel and bulk-copied (TMA engine) as a ready 16 KB hi|lo image per chunk
  const bool bimg = MODE != DW && p.b_img != nullptr;
  float va[kTcU][8], vb[kTcU][8];
  float bsum = 0.f;
  auto unit = [&](bool cmap, int q, int& r, int& g) {
    const int u = tid + q * kThreads;
    if (cmap) { r = ((u >> 5) << 3) | (u & 7); g = (u >> 3) & 3; } else { r = u & 127; g = u >> 7; }
  };
  auto fetch = [&](int n) {
    const int cb = c_begin + n * kTcK;
#pragma unroll
    for (int q = 0; q < kTcU; q++) {
      int r, g;
      unit(A_C, q, r, g);
      const int gi = i0 + r, c = cb + g * 8;
      if (gi >= p.I || c >= c_end) {
#pragma unroll
        for (int e = 0; e < 8; e++) va[q][e] = 0.f;
      } else if (MODE == FWD) {
        load8_rowsrc(p.x, gi, c, c_end, va[q]);
      } else if (MODE == DX) {
        load8(p.dP + (size_t)gi * p.ldp + c, c_end - c, va[q]);
      } else {
#pragma unroll
        for (int e = 0; e < 8; e++) va[q][e] = c + e < c_end ? __ldg(p.dP + (size_t)(c + e) * p.ldp + gi) : 0.f;
      }
    }
    if (!bimg)
#pragma unroll
    for (int q = 0; q < kTcU; q++) {
      int r, g;
      unit(B_C, q, r, g);
      const int gj = j0 + r, c = cb + g * 8;
      if (gj >= p.J || c >= c_end) {
#pragma unroll
        for (int e = 0; e < 8; e++) vb[q][e] = 0.f;
      } else if (MODE == FWD) {
        load8(p.W + (size_t)gj * p.ldw + c, c_end - c, vb[q]);
      } else if (MODE == DX) {
#pragma unroll
        for (int e = 0; e < 8; e++) vb[q][e] = c + e < c_end ? __ldg(p.W + (size_t)(c + e) * p.ldw + p.wk0 + gj) : 0.f;
      } else {
#pragma unroll
        for (int e = 0; e < 8; e++) vb[q][e] = c + e < c_end ? __ldg(p.x.ptr(c + e, gj)) : 0.f;
      }
    }
  };

  if (nchunks > 0) fetch(0);
  for (int n = 0; n < nchunks; n++) {
    const int st = n & 1;
    uint8_t* stage = tc_smem + st * kTcStage;
    if (n >= 2) mbar_wait(bar0 + 8 * st, ((n >> 1) - 1) & 1);      // MMAs of chunk n-2 have consumed this buffer
    if (bimg && tid == 0) {
      mbar_expect_tx(barf + 8 * st, kTcOp);
      bulk_g2s(smem_u32(stage + kTcOp), p.b_img + ((size_t)blockIdx.y * p.b_chunks + n) * kTcOp, kTcOp, barf + 8 * st);
    }
#pragma unroll
    for (int q = 0; q < kTcU; q++) {
      int r, g;
      unit(A_C, q, r, g);
      store_unit(stage, r, g, va[q], a_scale);
      if (MODE == DW && blockIdx.y == 0) {
#pragma unroll
        for (int e = 0; e < 8; e++) bsum += va[q][e];
      }
      if (!bimg) {
        unit(B_C, q, r, g);
        store_unit(stage + kTcOp, r, g, vb[q], b_scale);
      }
    }
    fence_proxy_async();
    if (n + 1 < nchunks) fetch(n + 1);                               // in flight across the barrier and the MMAs
    __syncthreads();
    if (tid == 0) {
      if (bimg) mbar_wait(barf + 8 * st, (n >> 1) & 1);
      fence_after_sync();
      const uint32_t a_hi = smem_u32(stage), a_lo = a_hi + kTcHalf, b_hi = a_hi + kTcOp, b_lo = b_hi + kTcHalf;
      const int rem = c_end - (c_begin + n * kTcK);
      const int nsl = rem >= kTcK ? kTcK / 16 : (rem + 15) / 16;
      for (int sl = 0; sl < nsl; sl++) {
        const uint32_t o = sl * 4096;
        mma_ss(tmem, smem_desc(a_hi + o, 2048, 128), smem_desc(b_hi + o, 2048, 128), idesc, (n > 0 || sl > 0) ? 1u : 0u);
        mma_ss(tmem, smem_desc(a_lo + o, 2048, 128), smem_desc(b_hi + o, 2048, 128), idesc, 1u);
        mma_ss(tmem, smem_desc(a_hi + o, 2048, 128), smem_desc(b_lo + o, 2048, 128), idesc, 1u);
      }
      mma_commit(bar0 + 8 * st);
    }
  }
  if (MODE == DW && blockIdx.y == 0) sbsum[tid] = bsum;               // row r = tid & 127, two threads per row
  if (nchunks > 0) {
    const int last = nchunks - 1;
    mbar_wait(bar0 + 8 * (last & 1), (last >> 1) & 1);
  }
  fence_after_sync();
  __syncthreads();

  // epilogue, phase 1: TMEM (lane = output row; warp w: lane quadrant w & 3, column half w >> 2) -> fp32 tile in
  // the (now idle) operand buffers, row stride 132 floats (conflict-free 16 B stores)
  constexpr int kLd = 132;
  float* tile = reinterpret_cast<float*>(tc_smem);
  const int q4 = warp & 3, hf = warp >> 2, row = q4 * 32 + lane;
  const uint32_t lane_base = tmem + ((uint32_t)(q4 * 32) << 16);
  const float inv = 1.f / (a_scale * b_scale);
  if (nchunks > 0 && (nb == 128 || hf == 0)) {
#pragma unroll 1
    for (int cc = 0; cc < 4; cc++) {
      const int col = hf * 64 + cc * 16;
      uint32_t u[16];
      tmem_ld16(lane_base + col, u);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 16; e += 4)
        *reinterpret_cast<float4*>(tile + row * kLd + col + e) =
            make_float4(__uint_as_float(u[e]) * inv, __uint_as_float(u[e + 1]) * inv, __uint_as_float(u[e + 2]) * inv,
                        __uint_as_float(u[e + 3]) * inv);
    }
  }
  __syncthreads();
  // phase 2: lanes along the output columns -> coalesced reads of bias / saved activations / addend and writes
  float* out = p.out;
  if (MODE == DW) out += (size_t)blockIdx.z * p.split_stride;
  float vmax = 0.f;
  const int ncol = nb;
  const bool vec = (p.ldo & 3) == 0 && (p.J & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                   (MODE != FWD || (reinterpret_cast<uintptr_t>(p.b) & 15) == 0) &&
                   (MODE != DX || ((!p.hprev || ((p.ldh & 3) == 0 && (reinterpret_cast<uintptr_t>(p.hprev) & 15) == 0)) &&
                                   !p.addend));
  if (vec) {
    // float4 per lane, four independent (row, column-quad) items per thread in flight: the saved-activation reads
    // of the DX epilogue are issued together instead of one dependent load per element
    const int q4n = ncol >> 2, total = 128 * q4n, sh = nb == 128 ? 5 : 4;
    constexpr int kB = 4;
    for (int base = tid; base < total; base += kThreads * kB) {
      float4 h[kB];
      int gi_[kB], gj_[kB];
      bool ok[kB];
#pragma unroll
      for (int bb = 0; bb < kB; bb++) {
        const int idx = base + bb * kThreads;
        const int r = idx >> sh, c4 = (idx & (q4n - 1)) << 2;
        gi_[bb] = i0 + r; gj_[bb] = j0 + c4;
        ok[bb] = idx < total && gi_[bb] < p.I && gj_[bb] < p.J;
        h[bb] = make_float4(1.f, 1.f, 1.f, 1.f);
        if (ok[bb]) {
          if (MODE == DX && p.hprev) h[bb] = __ldg(reinterpret_cast<const float4*>(p.hprev + (size_t)gi_[bb] * p.ldh + gj_[bb]));
          if (MODE == FWD) h[bb] = __ldg(reinterpret_cast<const float4*>(p.b + gj_[bb]));
        }
      }
#pragma unroll
      for (int bb = 0; bb < kB; bb++) {
        if (!ok[bb]) continue;
        const int idx = base + bb * kThreads;
        const int r = idx >> sh, c4 = (idx & (q4n - 1)) << 2;
        float4 v = nchunks > 0 ? *reinterpret_cast<const float4*>(tile + r * kLd + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE == FWD) {
          v.x = act_fwd(v.x + h[bb].x, p.act); v.y = act_fwd(v.y + h[bb].y, p.act);
          v.z = act_fwd(v.z + h[bb].z, p.act); v.w = act_fwd(v.w + h[bb].w, p.act);
        } else if (MODE == DX) {
          if (p.hprev) {
            v.x *= act_bwd(h[bb].x, p.act_prev); v.y *= act_bwd(h[bb].y, p.act_prev);
            v.z *= act_bwd(h[bb].z, p.act_prev); v.w *= act_bwd(h[bb].w, p.act_prev);
          }
          vmax = fmaxf(vmax, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        }
        *reinterpret_cast<float4*>(out + (size_t)gi_[bb] * p.ldo + gj_[bb]) = v;
      }
    }
  } else {
    for (int idx = tid; idx < 128 * ncol; idx += kThreads) {
      const int r = nb == 128 ? idx >> 7 : idx >> 6, cidx = idx & (ncol - 1);
      const int gi = i0 + r, gj = j0 + cidx;
      if (gi >= p.I || gj >= p.J) continue;
      float v = nchunks > 0 ? tile[r * kLd + cidx] : 0.f;
      if (MODE == FWD) {
        v = act_fwd(v + __ldg(p.b + gj), p.act);
      } else if (MODE == DX) {
        if (p.hprev) v *= act_bwd(__ldg(p.hprev + (size_t)gi * p.ldh + gj), p.act_prev);
        if (p.addend) v = fmaf(p.add_scale, __ldg(p.addend + (size_t)gi * p.J + gj), v);
        vmax = fmaxf(vmax, fabsf(v));
      }
      out[(size_t)gi * p.ldo + gj] = v;
    }
  }
  if (MODE == DX && p.amax_out) {
    for (int o = 16; o; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if (lane == 0 && vmax > 0.f) atomicMax(p.amax_out, __float_as_uint(vmax));
  }
  if (MODE == DW && blockIdx.y == 0 && tid < 128 && i0 + tid < p.I)
    p.db_out[(size_t)blockIdx.z * p.split_stride + i0 + tid] = sbsum[tid] + sbsum[tid + 128];
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc128(tmem);
}

template <int MODE>
static int launch_gemm_tc(const GemmP& p, int splits, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_devs{0};
  if (first_use_on_device(attr_devs)) {
    NDP_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem));
  }
  const dim3 grid((p.I + 127) / 128, (p.J + 127) / 128, splits);
  gemm_tc_kernel<MODE><<<grid, kThreads, kTcSmem, s>>>(p);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---- thin layers at large batch (fc5: 256 -> 4 actions, the logit: 256 -> 1): bandwidth-bound row sweeps -----------
// FWD, J <= 8: one warp per row, lanes along the contraction, W in shared memory
__global__ void __launch_bounds__(256) thin_fwd_kernel(const GemmP p) {
  extern __shared__ float thin_w[];                 // [J][Cn]
  const int J = p.J, K = p.Cn;
  for (int e = threadIdx.x; e < J * K; e += 256) thin_w[e] = __ldg(p.W + (size_t)(e / K) * p.ldw + e % K);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int m = blockIdx.x * 8 + warp; m < p.I; m += gridDim.x * 8) {
    const float* xr = p.x.p0 + (size_t)m * p.x.ld0;
    float acc[8];
#pragma unroll
    for (int n = 0; n < 8; n++) acc[n] = 0.f;
    for (int k = lane; k < K; k += 32) {
      const float xv = __ldg(xr + k);
#pragma unroll
      for (int n = 0; n < 8; n++)
        if (n < J) acc[n] = fmaf(xv, thin_w[n * K + k], acc[n]);
    }
#pragma unroll
    for (int n = 0; n < 8; n++) {
      if (n >= J) break;
      float v = acc[n];
      for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) p.out[(size_t)m * p.ldo + n] = act_fwd(v + __ldg(p.b + n), p.act);
    }
  }
}
// DW, I <= 8 output features: a warp sweeps rows of the split (8 warps -> 8 rows in flight, coalesced reads of X, the
// I gradient values of the row broadcast), lanes own 8 of the block's 256 input columns; the warps' partial sums are
// added in warp order (deterministic)
__global__ void __launch_bounds__(256) thin_dw_kernel(const GemmP p) {
  __shared__ float tile[8][256];
  __shared__ float sb[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * 256;
  const int c_begin = blockIdx.z * p.rows_per_split, c_end = min(p.Cn, c_begin + p.rows_per_split);
  float acc[8][8], bs[8];
#pragma unroll
  for (int n = 0; n < 8; n++) {
    bs[n] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; j++) acc[n][j] = 0.f;
  }
#pragma unroll 2
  for (int m = c_begin + warp; m < c_end; m += 8) {
    const float* xr = p.x.p0 + (size_t)m * p.x.ld0 + k0;
    float xv[8], d[8];
#pragma unroll
    for (int j = 0; j < 8; j++) xv[j] = (k0 + lane + 32 * j) < p.J ? __ldg(xr + lane + 32 * j) : 0.f;
#pragma unroll
    for (int n = 0; n < 8; n++) d[n] = n < p.I ? __ldg(p.dP + (size_t)m * p.ldp + n) : 0.f;
#pragma unroll
    for (int n = 0; n < 8; n++) {
      bs[n] += d[n];
#pragma unroll
      for (int j = 0; j < 8; j++) acc[n][j] = fmaf(d[n], xv[j], acc[n][j]);
    }
  }
  for (int w = 0; w < 8; w++) {
    if (warp == w) {
#pragma unroll
      for (int n = 0; n < 8; n++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (w == 0) tile[n][lane + 32 * j] = acc[n][j]; else tile[n][lane + 32 * j] += acc[n][j];
        }
        if (lane == 0) { if (w == 0) sb[n] = bs[n]; else sb[n] += bs[n]; }
      }
    }
    __syncthreads();
  }
  float* out = p.out + (size_t)blockIdx.z * p.split_stride;
  for (int e = threadIdx.x; e < p.I * 256; e += 256) {
    const int n = e >> 8, k = k0 + (e & 255);
    if (k < p.J) out[(size_t)n * p.ldo + k] = tile[n][e & 255];
  }
  if (blockIdx.x == 0 && threadIdx.x < p.I) p.db_out[(size_t)blockIdx.z * p.split_stride + threadIdx.x] = sb[threadIdx.x];
}

// 64-row tiles when they fill the GPU, 16-row tiles (4x the blocks, a quarter of the serial work each) otherwise:
// at the reference's batch (336 rows) a layer is a handful of 64x64 tiles and the step is latency-bound.
template <int MODE>
static int launch_gemm(const GemmP& p, int splits, cudaStream_t s) {
  // tensor path: wide layers at large batch (thin layers -- fc5, the logit, the 4 action columns -- stay on FFMA)
  const bool tc = p.use_tc && p.J >= 64 && p.Cn >= 64 &&
                  (MODE == DW ? (p.I >= 64 && p.Cn >= 2048 && p.rows_per_split % kTcK == 0) : p.I >= 2048);
  if (tc) return launch_gemm_tc<MODE>(p, splits, s);
  const bool plain_x = p.x.k1 == 0 && p.x.div0 == 1;
  if (MODE == FWD && p.J <= 8 && p.I >= 2048 && plain_x && (size_t)p.J * p.Cn * 4 <= 40960) {
    const int blocks = (p.I + 7) / 8 < 8 * num_sms_cached() ? (p.I + 7) / 8 : 8 * num_sms_cached();
    thin_fwd_kernel<<<blocks, 256, (size_t)p.J * p.Cn * 4, s>>>(p);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  if (MODE == DW && p.I <= 8 && p.Cn >= 2048 && plain_x) {
    thin_dw_kernel<<<dim3((p.J + 255) / 256, 1, splits), 256, 0, s>>>(p);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  const dim3 block(kThreads);
  const int tj = p.J >= 64 ? 64 : 16;
  const long long big_tiles = (long long)((p.I + 63) / 64) * ((p.J + tj - 1) / tj) * splits;
  const bool small = big_tiles < num_sms_cached() && (MODE != DW || p.Cn <= 4096);
  const int ti = small ? 16 : 64;
  const dim3 grid((p.I + ti - 1) / ti, (p.J + tj - 1) / tj, splits);
  if (tj == 64) {
    if (small) gemm_kernel<MODE, 16, 64><<<grid, block, 0, s>>>(p);
    else gemm_kernel<MODE, 64, 64><<<grid, block, 0, s>>>(p);
  } else {
    if (small) gemm_kernel<MODE, 16, 16><<<grid, block, 0, s>>>(p);
    else gemm_kernel<MODE, 64, 16><<<grid, block, 0, s>>>(p);
  }
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---- BCE-with-logits (mean) for up to two (logits, target) sets; writes the summed loss and d loss / d logit.
// nn.BCEWithLogitsLoss: l = (1-y) x + max(-x,0) + log(exp(-max(-x,0)) + exp(-x-max(-x,0))) = (1-y) x + softplus(-x).
// One block: the reduction order is fixed.  Also advances the Adam step counter of the optimiser that consumes
// these gradients (each optimiser step in train_gan.py is preceded by exactly one loss evaluation).
constexpr int kBceBlocks = 64;
__global__ void __launch_bounds__(1024) bce_kernel(const float* __restrict__ x0, int n0, float y0,
                                                    float* __restrict__ dl0, const float* __restrict__ x1, int n1,
                                                    float y1, float* __restrict__ dl1, float* __restrict__ loss_out,
                                                    int* __restrict__ step_counter, float world,
                                                    float* __restrict__ partial, unsigned* __restrict__ ticket) {
  __shared__ float red[32];
  __shared__ bool last;
  float tot = 0.f;
  for (int set = 0; set < 2; set++) {
    const float* x = set ? x1 : x0;
    const int n = set ? n1 : n0;
    const float y = set ? y1 : y0;
    float* dl = set ? dl1 : dl0;
    if (n <= 0) continue;
    const float inv = 1.f / ((float)n * world);   // data parallel: mean over the GLOBAL batch (equal shards)
    float s = 0.f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
      const float v = x[i];
      const float e = expf(-fabsf(v));
      s += (1.f - y) * v + fmaxf(-v, 0.f) + log1pf(e);
      const float sig = v >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
      dl[i] = (sig - y) * inv;
    }
    tot += s * inv;
  }
  for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = tot;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) {
      partial[blockIdx.x] = v;
      __threadfence();
      last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {           // fixed-order sum of the block partials: deterministic
    __threadfence();
    float v = 0.f;
    for (int i = 0; i < (int)gridDim.x; i++) v += __ldcg(partial + i);
    *loss_out = v;
    *step_counter += 1;
    *ticket = 0;
  }
}

// ---- Adam (torch.optim.Adam defaults: no weight decay, no amsgrad; train_gan.py:98-104) over one flat parameter
// buffer; the gradient is the fixed-order sum of the DW partial slots.  hp = {lr, beta1, beta2, eps} in double,
// as the Python floats the reference passes.
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, float* __restrict__ m,
                                                   float* __restrict__ v, const float* __restrict__ part,
                                                   int nparts, size_t stride, int P, const double* __restrict__ hp,
                                                   const int* __restrict__ step, float* __restrict__ grad_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  float g = 0.f;
  for (int s = 0; s < nparts; s++) g += part[(size_t)s * stride + i];
  if (grad_out) grad_out[i] = g;
  const double lr = hp[0], b1 = hp[1], b2 = hp[2], eps = hp[3];
  const int t = *step;
  const double bc1 = 1.0 - pow(b1, (double)t), bc2 = 1.0 - pow(b2, (double)t);
  const float step_size = (float)(lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  const float w1 = (float)(1.0 - b1);
  float mi = m[i], vi = v[i];
  mi = w1 < 0.5f ? mi + w1 * (g - mi) : g - (g - mi) * (1.f - w1);     // exp_avg.lerp_(grad, 1 - beta1)
  vi = vi * (float)b2 + (float)(1.0 - b2) * g * g;                      // mul_(beta2).addcmul_(grad, grad, 1-beta2)
  const float denom = sqrtf(vi) / bc2_sqrt + (float)eps;
  p[i] = p[i] - step_size * (mi / denom);
  m[i] = mi;
  v[i] = vi;
}

// data-parallel step: fixed-order sum of the partial slots into the flat gradient (then summed over ranks by NCCL)
__global__ void __launch_bounds__(256) reduce_parts_kernel(const float* __restrict__ part, int nparts, size_t stride,
                                                           int P, float* __restrict__ grad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  float g = 0.f;
  for (int s = 0; s < nparts; s++) g += part[(size_t)s * stride + i];
  grad[i] = g;
}

// ---- network description: flat parameter buffer [W0 | b0 | W1 | b1 | ...], torch [out,in] weights; weight images
// of the tensor path: per layer a FWD image (operand rows = out features, contraction = in) and a DX image (rows =
// in features, contraction = out), each [row tile][chunk][hi 8 KB | lo 8 KB], values x 2^8
struct Net {
  int L;
  int in[5], out[5];
  size_t w_off[5], b_off[5];
  size_t P;
  size_t fwd_off[5], dx_off[5];     // byte offsets into img; (size_t)-1 = layer stays on the exact kernel
  int fwd_chunks[5], dx_chunks[5];
  size_t img_bytes;
  uint8_t* img;
  void init(int L_, const int* in_, const int* out_) {
    L = L_;
    P = 0;
    img_bytes = 0;
    img = nullptr;
    for (int l = 0; l < L; l++) {
      in[l] = in_[l]; out[l] = out_[l];
      w_off[l] = P; P += (size_t)in[l] * out[l];
      b_off[l] = P; P += out[l];
      const bool wide = in[l] >= 64 && out[l] >= 64;
      fwd_chunks[l] = (in[l] + kTcK - 1) / kTcK;
      dx_chunks[l] = (out[l] + kTcK - 1) / kTcK;
      fwd_off[l] = dx_off[l] = (size_t)-1;
      if (wide) {
        fwd_off[l] = img_bytes; img_bytes += (size_t)((out[l] + 127) / 128) * fwd_chunks[l] * kTcOp;
        dx_off[l] = img_bytes; img_bytes += (size_t)((in[l] + 127) / 128) * dx_chunks[l] * kTcOp;
      }
    }
  }
};

struct PackDesc {
  int L;
  int in[5], out[5];
  size_t w_off[5], fwd_off[5], dx_off[5];
};
// grid (units / 256, 2 * L): blockIdx.y = 2 * layer + kind (0 FWD image, 1 DX image); one (row, 8-group) unit per
// thread, lanes along the operand rows
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ params, uint8_t* __restrict__ img,
                                                           const PackDesc d) {
  const int l = blockIdx.y >> 1, kind = blockIdx.y & 1;
  const size_t off = kind ? d.dx_off[l] : d.fwd_off[l];
  if (off == (size_t)-1) return;
  const int R = kind ? d.in[l] : d.out[l], K = kind ? d.out[l] : d.in[l];
  const int chunks = (K + kTcK - 1) / kTcK, rtiles = (R + 127) / 128;
  const int u = blockIdx.x * 256 + threadIdx.x;
  const int per_tile = 128 * (kTcK / 8);
  if (u >= rtiles * chunks * per_tile) return;
  const int tile = u / per_tile, w = u % per_tile;
  const int r = w & 127, g = w >> 7;
  const int rt = tile / chunks, ch = tile % chunks;
  const int gr = rt * 128 + r, k0 = ch * kTcK + g * 8;
  const float* W = params + d.w_off[l];
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; e++) {
    const int k = k0 + e;
    v[e] = (gr < R && k < K) ? (kind ? W[(size_t)k * d.in[l] + gr] : W[(size_t)gr * d.in[l] + k]) : 0.f;
  }
  store_unit(img + off + (size_t)tile * kTcOp, r, g, v, kUnitScale);
}
static int pack_weights(const Net& n, const float* params, cudaStream_t s) {
  if (!n.img || !n.img_bytes) return NDP_OK;
  PackDesc d;
  d.L = n.L;
  int max_units = 0;
  for (int l = 0; l < n.L; l++) {
    d.in[l] = n.in[l]; d.out[l] = n.out[l]; d.w_off[l] = n.w_off[l]; d.fwd_off[l] = n.fwd_off[l]; d.dx_off[l] = n.dx_off[l];
    const int uf = ((n.out[l] + 127) / 128) * n.fwd_chunks[l] * 128 * (kTcK / 8);
    const int ud = ((n.in[l] + 127) / 128) * n.dx_chunks[l] * 128 * (kTcK / 8);
    max_units = max_units > uf ? max_units : uf;
    max_units = max_units > ud ? max_units : ud;
  }
  pack_weights_kernel<<<dim3((max_units + 255) / 256, 2 * n.L), 256, 0, s>>>(params, n.img, d);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

constexpr int kMaxSplits = 128;
// batch rows -> partial slots of the weight-gradient products: 64-row slots while the batch is small (the step is
// latency-bound there), >= 512-row slots at large batch so that (output tiles x splits) fills the GPU twice over
static void split_plan(int M, int* splits, int* rows_per_split) {
  int s = M <= 4096 ? (M + 63) / 64 : (M + 511) / 512;
  const int cap = M <= 4096 ? 32 : kMaxSplits;
  if (s > cap) s = cap;
  if (M > 4096 && s < 32) s = 32;
  if (s < 1) s = 1;
  int rps = (M + s - 1) / s;
  rps = (rps + 63) / 64 * 64;
  *rows_per_split = rps;
  *splits = (M + rps - 1) / rps;
}

// acts[l] = output of layer l ([M x out[l]]); hidden activation `act`, none on the last layer
static int mlp_forward(const Net& n, const float* params, const RowSrc& x0, int M, float* const* acts, int act,
                       int use_tc, cudaStream_t s) {
  for (int l = 0; l < n.L; l++) {
    GemmP p = {};
    p.I = M; p.J = n.out[l]; p.Cn = n.in[l];
    p.x = l == 0 ? x0 : row_src(acts[l - 1], n.in[l], 1);
    p.W = params + n.w_off[l]; p.ldw = n.in[l];
    p.b = params + n.b_off[l];
    p.act = l + 1 < n.L ? act : ACT_NONE;
    p.out = acts[l]; p.ldo = n.out[l];
    p.use_tc = use_tc;
    if (n.img && n.fwd_off[l] != (size_t)-1) { p.b_img = n.img + n.fwd_off[l]; p.b_chunks = n.fwd_chunks[l]; }
    int rc = launch_gemm<FWD>(p, 1, s);
    if (rc) return rc;
  }
  return NDP_OK;
}

// event pool for the fork/join edges between the step's streams (record + wait pairs; reused round-robin)
struct EvPool {
  cudaEvent_t e[64];
  int n;
};
static int hop(EvPool& ev, cudaStream_t from, cudaStream_t to) {   // `to` continues after everything queued on `from`
  if (from == to) return NDP_OK;
  cudaEvent_t e = ev.e[ev.n++ & 63];
  NDP_CUDA(cudaEventRecord(e, from));
  NDP_CUDA(cudaStreamWaitEvent(to, e, 0));
  return NDP_OK;
}

// d_out = d loss / d (last layer output) [M x out[L-1]].  part != NULL: weight/bias gradient partials into slots
// [slot0, slot0+splits) of part (slot stride n.P).  dx0 != NULL: d loss / d input columns [col0, col0+ncols) of
// layer 0 (+ add_scale * addend).  dbuf[l]: d loss / d (output of layer l), l < L-1 (one buffer per layer, so the
// weight-gradient products may trail the dX chain on another stream).  The dX chain runs on s, the dW products on
// s_dw (== s: everything in order on one stream); the caller joins s_dw before it consumes the partials.
static int mlp_backward(const Net& n, const float* params, const RowSrc& x0, int M, float* const* acts,
                        const float* d_out, float* const* dbuf, int act, float* part, int slot0,
                        float* dx0, int col0, int ncols, const float* addend, float add_scale, int use_tc,
                        unsigned* amax, int* amax_used, cudaStream_t s, cudaStream_t s_dw, EvPool& ev) {
  int splits, rps;
  split_plan(M, &splits, &rps);
  const unsigned* cur_amax = nullptr;     // d_out feeds thin (FFMA) products only; every dP made here gets a slot
  const bool side = part && s_dw != s;
  if (side)
    if (int rc = hop(ev, s, s_dw)) return rc;       // d_out and the saved activations are ready
  for (int l = n.L - 1; l >= 0; l--) {
    const float* cur = l == n.L - 1 ? d_out : dbuf[l];
    if (part) {
      GemmP p = {};
      p.I = n.out[l]; p.J = n.in[l]; p.Cn = M;
      p.dP = cur; p.ldp = n.out[l];
      p.x = l == 0 ? x0 : row_src(acts[l - 1], n.in[l], 1);
      p.out = part + (size_t)slot0 * n.P + n.w_off[l]; p.ldo = n.in[l];
      p.db_out = part + (size_t)slot0 * n.P + n.b_off[l];
      p.rows_per_split = rps; p.split_stride = n.P;
      p.use_tc = use_tc; p.amax_in = cur_amax;
      int rc = launch_gemm<DW>(p, splits, s_dw);
      if (rc) return rc;
    }
    if (l > 0 || dx0) {
      GemmP p = {};
      p.I = M; p.Cn = n.out[l];
      p.dP = cur; p.ldp = n.out[l];
      p.W = params + n.w_off[l]; p.ldw = n.in[l];
      p.use_tc = use_tc; p.amax_in = cur_amax;
      unsigned* slot = nullptr;
      if (l > 0) {
        slot = amax + (*amax_used)++;
        p.amax_out = slot;
        p.J = n.in[l]; p.wk0 = 0;
        if (n.img && n.dx_off[l] != (size_t)-1) { p.b_img = n.img + n.dx_off[l]; p.b_chunks = n.dx_chunks[l]; }
        p.hprev = acts[l - 1]; p.ldh = n.in[l]; p.act_prev = act;
        p.out = dbuf[l - 1]; p.ldo = n.in[l];
      } else {
        p.J = ncols; p.wk0 = col0;
        p.addend = addend; p.add_scale = add_scale;
        p.out = dx0; p.ldo = ncols;
      }
      int rc = launch_gemm<DX>(p, 1, s);
      if (rc) return rc;
      cur_amax = slot;
      if (side && l > 0)
        if (int rc2 = hop(ev, s, s_dw)) return rc2;   // dP of layer l-1 is ready for its dW product
    }
  }
  return NDP_OK;
}

}  // namespace gan
}  // namespace ndp

// ---- handle ------------------------------------------------------------------------------------------------
using namespace ndp;
using namespace ndp::gan;

struct ndp_gan {
  ndp_gan_desc d;
  Net G, D;
  int B_max;
  void* arena;
  float *gp, *gm, *gv, *gg, *dp, *dm, *dv, *dg;
  int use_tc;        // tcgen05 products for wide layers at large batch (NDP_GAN_EXACT=1 keeps everything on FFMA)
  unsigned* amax;    // running max |dP| slots of one step (operand scales of the tensor path)
  int n_amax;
  double* hp;        // {lr, beta1, beta2, eps}
  int* steps;        // [0] discriminator, [1] generator optimiser step counts
  float* g_act[5];   // generator layer outputs, B rows; g_act[4] = a_hat
  float* dr_act[4];  // discriminator on the F distinct real rows
  float* df_act[4];  // discriminator on the B generated rows
  float *dl_real, *dl_fake, *g_div, *da, *sx, *sz;
  float* gd[4];      // d loss / d (generator layer l output), B rows
  float* fd[3];      // the same for the discriminator on the generated rows
  float* rd[3];      // ... and on the real rows (F rows)
  cudaStream_t s_real, s_dw, s_nd;   // side streams of the small-batch schedule (real branch, dW products, ndiv)
  EvPool ev;
  int multi_stream;
  float* bce_part;       // [kBceBlocks] block partials of the loss mean
  unsigned* bce_ticket;  // last-block ticket (self-resetting)
  float *g_part, *d_part;
  void* ndiv_ws;
  size_t ndiv_ws_bytes;
};

__global__ void set_hp_kernel(double* hp, double lr, double b1, double b2, double eps, int mask) {
  if (mask & 1) hp[0] = lr;
  if (mask & 2) hp[1] = b1;
  if (mask & 4) hp[2] = b2;
  if (mask & 8) hp[3] = eps;
}

static size_t carve(size_t& off, size_t bytes) {
  const size_t o = off;
  off += (bytes + 255) & ~(size_t)255;
  return o;
}

extern "C" {

int ndp_gan_create(ndp_gan** out, const ndp_gan_desc* d) {
  NDP_CHECK_ARG(out && d, "ndp_gan_create: null argument");
  int rc = device_check();
  if (rc) return rc;
  NDP_CHECK_ARG(d->C > 0 && d->nd > 0 && d->A > 0 && d->S > 0 && d->max_F > 0, "ndp_gan_create: bad sizes");
  NDP_CHECK_ARG(d->d_steps >= 0, "ndp_gan_create: d_steps < 0");
  NDP_CHECK_ARG((long long)d->max_F * d->S < (1ll << 24), "ndp_gan_create: max_F * S must stay below 2^24 rows");
  ndp_gan* g = new ndp_gan();
  g->d = *d;
  const int gin[5] = {d->C + d->nd, 128, 64, 128, 256}, gout[5] = {128, 64, 128, 256, d->A};   // gan.py:67-71
  const int din[4] = {d->A + d->C, 64, 128, 256}, dout[4] = {64, 128, 256, 1};                 // gan.py:93-96
  g->G.init(5, gin, gout);
  g->D.init(4, din, dout);
  const size_t F = d->max_F, B = F * d->S;
  g->B_max = (int)B;
  size_t ndiv_bytes = 0;
  rc = ndp_ndiv_workspace_bytes(d->max_F, d->S, d->A, d->nd, &ndiv_bytes);
  if (rc) { delete g; return rc; }
  g->ndiv_ws_bytes = ndiv_bytes;

  size_t off = 0;
  const size_t PG = g->G.P, PD = g->D.P;
  const size_t o_gp = carve(off, 4 * PG * 4), o_dp = carve(off, 4 * PD * 4);
  const size_t o_hp = carve(off, 4 * sizeof(double)), o_steps = carve(off, 2 * sizeof(int));
  g->n_amax = 8 * d->d_steps + 16;
  const size_t o_amax = carve(off, (size_t)g->n_amax * sizeof(unsigned));
  const char* ex = getenv("NDP_GAN_EXACT");
  g->use_tc = !(ex && ex[0] == '1');
  size_t o_g[5], o_dr[4], o_df[4];
  for (int l = 0; l < 5; l++) o_g[l] = carve(off, B * gout[l] * 4);
  for (int l = 0; l < 4; l++) o_dr[l] = carve(off, F * dout[l] * 4);
  for (int l = 0; l < 4; l++) o_df[l] = carve(off, B * dout[l] * 4);
  const size_t o_dlr = carve(off, F * 4), o_dlf = carve(off, B * 4);
  const size_t o_bcep = carve(off, kBceBlocks * 4), o_bcet = carve(off, 4);
  size_t o_gd[4], o_fd[3], o_rd[3];
  for (int l = 0; l < 4; l++) o_gd[l] = carve(off, B * gout[l] * 4);
  for (int l = 0; l < 3; l++) { o_fd[l] = carve(off, B * dout[l] * 4); o_rd[l] = carve(off, F * dout[l] * 4); }
  const size_t o_gdiv = carve(off, B * d->A * 4), o_da = carve(off, B * d->A * 4);
  const size_t o_sx = carve(off, B * 4), o_sz = carve(off, B * 4);
  const size_t o_gpart = carve(off, (size_t)kMaxSplits * PG * 4), o_dpart = carve(off, (size_t)2 * kMaxSplits * PD * 4);
  const size_t o_ws = carve(off, ndiv_bytes);
  const size_t o_gimg = carve(off, g->G.img_bytes), o_dimg = carve(off, g->D.img_bytes);
  cudaError_t e = cudaMalloc(&g->arena, off);
  if (e != cudaSuccess) {
    set_error("ndp_gan_create: cudaMalloc(%zu) failed: %s", off, cudaGetErrorString(e));
    delete g;
    return NDP_ERR_CUDA;
  }
  char* base = (char*)g->arena;
  g->gp = (float*)(base + o_gp); g->gm = g->gp + PG; g->gv = g->gm + PG; g->gg = g->gv + PG;
  g->dp = (float*)(base + o_dp); g->dm = g->dp + PD; g->dv = g->dm + PD; g->dg = g->dv + PD;
  g->hp = (double*)(base + o_hp); g->steps = (int*)(base + o_steps);
  g->amax = (unsigned*)(base + o_amax);
  for (int l = 0; l < 5; l++) g->g_act[l] = (float*)(base + o_g[l]);
  for (int l = 0; l < 4; l++) { g->dr_act[l] = (float*)(base + o_dr[l]); g->df_act[l] = (float*)(base + o_df[l]); }
  g->dl_real = (float*)(base + o_dlr); g->dl_fake = (float*)(base + o_dlf);
  g->bce_part = (float*)(base + o_bcep); g->bce_ticket = (unsigned*)(base + o_bcet);
  NDP_CUDA(cudaMemset(g->bce_ticket, 0, 4));
  for (int l = 0; l < 4; l++) g->gd[l] = (float*)(base + o_gd[l]);
  for (int l = 0; l < 3; l++) { g->fd[l] = (float*)(base + o_fd[l]); g->rd[l] = (float*)(base + o_rd[l]); }
  const char* ss = getenv("NDP_GAN_SINGLE_STREAM");
  g->multi_stream = !(ss && ss[0] == '1');
  bool ok = cudaStreamCreateWithFlags(&g->s_real, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&g->s_dw, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&g->s_nd, cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; i < 64 && ok; i++) ok = cudaEventCreateWithFlags(&g->ev.e[i], cudaEventDisableTiming) == cudaSuccess;
  g->ev.n = 0;
  if (!ok) {
    set_error("ndp_gan_create: stream / event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
    cudaFree(g->arena);
    delete g;
    return NDP_ERR_CUDA;
  }
  g->g_div = (float*)(base + o_gdiv); g->da = (float*)(base + o_da);
  g->sx = (float*)(base + o_sx); g->sz = (float*)(base + o_sz);
  g->g_part = (float*)(base + o_gpart); g->d_part = (float*)(base + o_dpart);
  g->ndiv_ws = base + o_ws;
  if (g->use_tc) {
    g->G.img = (uint8_t*)(base + o_gimg);
    g->D.img = (uint8_t*)(base + o_dimg);
  }
  *out = g;
  return NDP_OK;
}

void ndp_gan_destroy(ndp_gan* g) {
  if (!g) return;
  cudaStreamDestroy(g->s_real); cudaStreamDestroy(g->s_dw); cudaStreamDestroy(g->s_nd);
  for (int i = 0; i < 64; i++) cudaEventDestroy(g->ev.e[i]);
  cudaFree(g->arena);
  delete g;
}

// Copies the parameters (torch [out,in] weights, device pointers) into the handle and resets both optimisers.
int ndp_gan_load(ndp_gan* g, const float* const* gW, const float* const* gb, const float* const* dW,
                 const float* const* db, void* stream) {
  NDP_CHECK_ARG(g && gW && gb && dW && db, "ndp_gan_load: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  for (int l = 0; l < g->G.L; l++) {
    NDP_CUDA(cudaMemcpyAsync(g->gp + g->G.w_off[l], gW[l], (size_t)g->G.in[l] * g->G.out[l] * 4, cudaMemcpyDeviceToDevice, s));
    NDP_CUDA(cudaMemcpyAsync(g->gp + g->G.b_off[l], gb[l], (size_t)g->G.out[l] * 4, cudaMemcpyDeviceToDevice, s));
  }
  for (int l = 0; l < g->D.L; l++) {
    NDP_CUDA(cudaMemcpyAsync(g->dp + g->D.w_off[l], dW[l], (size_t)g->D.in[l] * g->D.out[l] * 4, cudaMemcpyDeviceToDevice, s));
    NDP_CUDA(cudaMemcpyAsync(g->dp + g->D.b_off[l], db[l], (size_t)g->D.out[l] * 4, cudaMemcpyDeviceToDevice, s));
  }
  NDP_CUDA(cudaMemsetAsync(g->gm, 0, 3 * g->G.P * 4, s));
  NDP_CUDA(cudaMemsetAsync(g->dm, 0, 3 * g->D.P * 4, s));
  NDP_CUDA(cudaMemsetAsync(g->steps, 0, 2 * sizeof(int), s));
  set_hp_kernel<<<1, 1, 0, s>>>(g->hp, g->d.lr, g->d.beta1, g->d.beta2, g->d.eps, 15);
  NDP_LAUNCH_CHECK();
  if (int rc = pack_weights(g->G, g->gp, s)) return rc;
  if (int rc = pack_weights(g->D, g->dp, s)) return rc;
  return NDP_OK;
}

int ndp_gan_export(ndp_gan* g, float* const* gW, float* const* gb, float* const* dW, float* const* db,
                   void* stream) {
  NDP_CHECK_ARG(g && gW && gb && dW && db, "ndp_gan_export: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  for (int l = 0; l < g->G.L; l++) {
    NDP_CUDA(cudaMemcpyAsync(gW[l], g->gp + g->G.w_off[l], (size_t)g->G.in[l] * g->G.out[l] * 4, cudaMemcpyDeviceToDevice, s));
    NDP_CUDA(cudaMemcpyAsync(gb[l], g->gp + g->G.b_off[l], (size_t)g->G.out[l] * 4, cudaMemcpyDeviceToDevice, s));
  }
  for (int l = 0; l < g->D.L; l++) {
    NDP_CUDA(cudaMemcpyAsync(dW[l], g->dp + g->D.w_off[l], (size_t)g->D.in[l] * g->D.out[l] * 4, cudaMemcpyDeviceToDevice, s));
    NDP_CUDA(cudaMemcpyAsync(db[l], g->dp + g->D.b_off[l], (size_t)g->D.out[l] * 4, cudaMemcpyDeviceToDevice, s));
  }
  return NDP_OK;
}

// Gradients of the last step in the flat layout [W0 | b0 | W1 | b1 | ...]: generator (of G_loss + factor*div),
// discriminator (of D_loss, last discriminator step).  Either pointer may be NULL.
int ndp_gan_param_counts(const ndp_gan* g, size_t* n_gen, size_t* n_dis) {
  NDP_CHECK_ARG(g && n_gen && n_dis, "ndp_gan_param_counts: null argument");
  *n_gen = g->G.P;
  *n_dis = g->D.P;
  return NDP_OK;
}
int ndp_gan_grads(ndp_gan* g, float* gen_flat, float* dis_flat, void* stream) {
  NDP_CHECK_ARG(g, "ndp_gan_grads: null handle");
  cudaStream_t s = (cudaStream_t)stream;
  if (gen_flat) NDP_CUDA(cudaMemcpyAsync(gen_flat, g->gg, g->G.P * 4, cudaMemcpyDeviceToDevice, s));
  if (dis_flat) NDP_CUDA(cudaMemcpyAsync(dis_flat, g->dg, g->D.P * 4, cudaMemcpyDeviceToDevice, s));
  return NDP_OK;
}

int ndp_gan_set_lr(ndp_gan* g, double lr, void* stream) {
  NDP_CHECK_ARG(g && lr >= 0.0, "ndp_gan_set_lr: bad argument");
  set_hp_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(g->hp, lr, 0, 0, 0, 1);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// gradient = fixed-order sum of the partial slots; with a communicator, additionally summed over the ranks
// (ONE all-reduce per optimiser step, on the caller's stream) before the update
static int adam(float* p, float* m, float* v, float* grad, const float* part, int nparts, size_t P, const double* hp,
                const int* step, ndp_comm* comm, cudaStream_t s) {
  const unsigned blocks = (unsigned)((P + 255) / 256);
  if (comm) {
    reduce_parts_kernel<<<blocks, 256, 0, s>>>(part, nparts, P, (int)P, grad);
    NDP_LAUNCH_CHECK();
    if (int rc = ndp_comm_allreduce_sum_f32(comm, grad, (int)P, s)) return rc;
    adam_kernel<<<blocks, 256, 0, s>>>(p, m, v, grad, 1, 0, (int)P, hp, step, nullptr);
  } else {
    adam_kernel<<<blocks, 256, 0, s>>>(p, m, v, part, nparts, P, (int)P, hp, step, grad);
  }
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// One iteration of the loop body train_gan.py:159-203 for F (state, goal) codes with S samples each.
// codes [F,C], actions [F,A], noise [F,S,nd] (device, fp32); losses [3] (device) = {D_loss, G_loss, pair_div_loss}.
static int train_step_impl(ndp_gan* g, ndp_comm* comm, int world, const float* codes, const float* actions,
                           const float* noise, int F, float* losses, void* stream) {
  NDP_CHECK_ARG(g && codes && actions && noise && losses, "ndp_gan_train_step: null argument");
  NDP_CHECK_ARG(F > 0 && F <= g->d.max_F, "ndp_gan_train_step: F=%d outside (0, max_F=%d]", F, g->d.max_F);
  cudaStream_t s = (cudaStream_t)stream;
  const int S = g->d.S, C = g->d.C, A = g->d.A, nd = g->d.nd, B = F * S;
  int rc, used = 0;
  const int tc = g->use_tc;
  const int bce_blocks = B <= 1024 ? 1 : (B / 1024 < kBceBlocks ? B / 1024 : kBceBlocks);
  NDP_CUDA(cudaMemsetAsync(g->amax, 0, (size_t)g->n_amax * sizeof(unsigned), s));
#define NDP_TRY(x) do { rc = (x); if (rc) return rc; } while (0)
  // generator forward on [codes repeated S times | noise]                      (train_gan.py:159-165)
  const RowSrc xg = row_src(codes, C, S, noise, nd, 1);
  NDP_TRY(mlp_forward(g->G, g->gp, xg, B, g->g_act, ACT_RELU, tc, s));
  const float* a_hat = g->g_act[4];
  const RowSrc x_real = row_src(actions, A, 1, codes, C, 1);     // cat([action, state_code]) gan.py:102
  const RowSrc x_fake = row_src(a_hat, A, 1, codes, C, S);
  int sr, sf, rps;
  split_plan(F, &sr, &rps);
  split_plan(B, &sf, &rps);
  // Small batches are latency-bound (every kernel is a handful of blocks), so independent work runs side by side:
  // the real-row branch of the discriminator, the weight-gradient products (off the dX chain) and the ndiv loss each
  // get a side stream, forked from and joined into the caller's stream with events (capturable in a CUDA graph).
  // Large batches fill the GPU kernel by kernel and stay on one stream.
  const bool ms = g->multi_stream && B <= 8192;
  cudaStream_t s_real = ms ? g->s_real : s, s_dw = ms ? g->s_dw : s, s_nd = ms ? g->s_nd : s;
  EvPool& ev = g->ev;
  // normalized diversification of the S samples of every code                   (:193-196)
  NDP_TRY(hop(ev, s, s_nd));
  NDP_TRY(ndp_ndiv_forward_backward(a_hat, noise, F, S, A, nd, g->d.margin, g->sx, g->sz, losses + 2, g->g_div,
                                    g->ndiv_ws, g->ndiv_ws_bytes, (void*)s_nd));
  for (int it = 0; it < g->d.d_steps; it++) {                    // :172-184
    NDP_TRY(hop(ev, s, s_real));
    NDP_TRY(mlp_forward(g->D, g->dp, x_real, F, g->dr_act, ACT_LEAKY, tc, s_real));
    NDP_TRY(mlp_forward(g->D, g->dp, x_fake, B, g->df_act, ACT_LEAKY, tc, s));
    NDP_TRY(hop(ev, s_real, s));
    bce_kernel<<<bce_blocks, 1024, 0, s>>>(g->dr_act[3], F, 1.f, g->dl_real, g->df_act[3], B, 0.f, g->dl_fake,
                                           losses + 0, g->steps + 0, (float)world, g->bce_part, g->bce_ticket);
    NDP_LAUNCH_CHECK();
    NDP_TRY(hop(ev, s, s_real));
    NDP_TRY(mlp_backward(g->D, g->dp, x_real, F, g->dr_act, g->dl_real, g->rd, ACT_LEAKY, g->d_part, 0,
                         nullptr, 0, 0, nullptr, 0.f, tc, g->amax, &used, s_real, s_real, ev));
    NDP_TRY(mlp_backward(g->D, g->dp, x_fake, B, g->df_act, g->dl_fake, g->fd, ACT_LEAKY, g->d_part, sr,
                         nullptr, 0, 0, nullptr, 0.f, tc, g->amax, &used, s, s_dw, ev));
    NDP_TRY(hop(ev, s_real, s));
    NDP_TRY(hop(ev, s_dw, s));
    NDP_TRY(adam(g->dp, g->dm, g->dv, g->dg, g->d_part, sr + sf, g->D.P, g->hp, g->steps + 0, comm, s));
    NDP_TRY(pack_weights(g->D, g->dp, s));         // images always follow the parameters (F may change per call)
  }
  // generator loss through the UPDATED discriminator                            (:187-190)
  NDP_TRY(mlp_forward(g->D, g->dp, x_fake, B, g->df_act, ACT_LEAKY, tc, s));
  bce_kernel<<<bce_blocks, 1024, 0, s>>>(g->df_act[3], B, 1.f, g->dl_fake, nullptr, 0, 0.f, nullptr, losses + 1,
                                         g->steps + 1, (float)world, g->bce_part, g->bce_ticket);
  NDP_LAUNCH_CHECK();
  NDP_TRY(hop(ev, s_nd, s));                                     // d div / d a_hat is needed from here on
  // d total / d a_hat = dD/da_hat (first A input columns of discriminator fc1) + div_factor * d div / d a_hat
  NDP_TRY(mlp_backward(g->D, g->dp, x_fake, B, g->df_act, g->dl_fake, g->fd, ACT_LEAKY, nullptr, 0, g->da,
                       0, A, g->g_div, g->d.div_factor, tc, g->amax, &used, s, s, ev));
  NDP_TRY(mlp_backward(g->G, g->gp, xg, B, g->g_act, g->da, g->gd, ACT_RELU, g->g_part, 0, nullptr, 0, 0,
                       nullptr, 0.f, tc, g->amax, &used, s, s_dw, ev));
  NDP_TRY(hop(ev, s_dw, s));
  NDP_TRY(adam(g->gp, g->gm, g->gv, g->gg, g->g_part, sf, g->G.P, g->hp, g->steps + 1, comm, s));   // :198-203
  NDP_TRY(pack_weights(g->G, g->gp, s));
  if (comm) NDP_TRY(ndp_comm_allreduce_sum_f32(comm, losses, 3, s));   // partial means / partial ndiv sum -> global
#undef NDP_TRY
  return NDP_OK;
}

int ndp_gan_train_step(ndp_gan* g, const float* codes, const float* actions, const float* noise, int F,
                       float* losses, void* stream) {
  return train_step_impl(g, nullptr, 1, codes, actions, noise, F, losses, stream);
}

// Data-parallel step: every rank holds the same parameters and optimiser state and F of the world * F codes of the
// global batch (equal shards).  Local forward/backward, ONE NCCL all-reduce of the flat gradient per optimiser step
// (discriminator: 58,305 floats, generator: 83,780 + nd*128) and one of the three loss scalars; the ndiv groups
// are rank-local.  Parameters stay bit-identical across ranks (same reduced gradient, same update).
int ndp_gan_train_step_dp(ndp_gan* g, ndp_comm* comm, int world, const float* codes, const float* actions,
                          const float* noise, int F, float* losses, void* stream) {
  NDP_CHECK_ARG(comm && world >= 1, "ndp_gan_train_step_dp: needs a communicator and world >= 1");
  return train_step_impl(g, comm, world, codes, actions, noise, F, losses, stream);
}

// a_hat of the last step ([F*S, A]) — what train_gan.py:165 calls action_hat
int ndp_gan_last_actions(ndp_gan* g, float* a_hat, int F, void* stream) {
  NDP_CHECK_ARG(g && a_hat && F > 0 && F <= g->d.max_F, "ndp_gan_last_actions: bad argument");
  NDP_CUDA(cudaMemcpyAsync(a_hat, g->g_act[4], (size_t)F * g->d.S * g->d.A * 4, cudaMemcpyDeviceToDevice,
                           (cudaStream_t)stream));
  return NDP_OK;
}

}  // extern "C"
