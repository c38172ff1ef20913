// K2 — normalized diversification loss (reference diversity.py:8-41), forward + backward,
// without ever materialising a k x k matrix in HBM.
//
//   d_ij  = || v_i - v_j ||_2                      compute_pairwise        diversity.py:8-9
//   delta = d_ij / sum_j d_ij (normaliser detached) compute_pair_distance   diversity.py:12-19
//   L     = sum relu(0.8 * delta^z_ij - delta^x_ij) compute_pairwise_divergence :36-41
//   dL/dx_i = sum_j What_ij (x_i - x_j),  What_ij = -(M_ij/sx_i + M_ji/sx_j) / dx_ij  (0 if dx_ij==0)
//
// Arithmetic is the reference's own (direct differences, fp32, IEEE sqrt/div) so parity
// does not depend on a Gram-form cancellation.  Two passes over the pair tiles:
//   pass 1  row sums sx, sz                          (ndp_ndiv_rowsums)
//   pass 2  loss partials + What tile + What*X GEMM  (ndp_ndiv_loss_grad)
// A block owns 32 complete rows and sweeps all j in 128-wide tiles -> no atomics,
// deterministic sums.  Algorithmic HBM bytes: 4*N*k*(Dz+2*Dx) + 8*N*k.
#include "common.cuh"

namespace ndp {

static constexpr int kTI = 32;        // rows per block
static constexpr int kTJ = 128;       // cols per tile
static constexpr int kDC = 16;        // feature chunk staged in smem
static constexpr int kNdivThreads = 256;
static constexpr int kPadI = kTI + 4;
static constexpr int kPadJ = kTJ + 4;

// stage v[rows r0..r0+R) x dims d0..d0+kDC) transposed into s[d][r] (zero padded)
template <int R, int PAD>
__device__ __forceinline__ void stage_chunk(float* s, const float* __restrict__ v, int k, int D,
                                            int r0, int d0) {
  for (int t = threadIdx.x; t < R * kDC; t += kNdivThreads) {
    int r = t / kDC, d = t % kDC;
    float val = 0.f;
    if (r0 + r < k && d0 + d < D) val = __ldg(v + (size_t)(r0 + r) * D + d0 + d);
    s[d * PAD + r] = val;
  }
}

// accumulate squared distances of the 4x4 pair sub-tile owned by this thread
__device__ __forceinline__ void sqdist_tile(float (&acc)[4][4], const float* __restrict__ vg, int k, int D,
                                            int i0, int j0, float* s_i, float* s_j) {
  const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
  for (int d0 = 0; d0 < D; d0 += kDC) {
    __syncthreads();
    stage_chunk<kTI, kPadI>(s_i, vg, k, D, i0, d0);
    stage_chunk<kTJ, kPadJ>(s_j, vg, k, D, j0, d0);
    __syncthreads();
    const int dn = (D - d0) < kDC ? (D - d0) : kDC;
    for (int d = 0; d < dn; ++d) {
      const float4 vi = *reinterpret_cast<const float4*>(s_i + d * kPadI + ty * 4);
      const float4 vj = *reinterpret_cast<const float4*>(s_j + d * kPadJ + tx * 4);
      const float ai[4] = {vi.x, vi.y, vi.z, vi.w};
      const float bj[4] = {vj.x, vj.y, vj.z, vj.w};
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const float df = ai[a] - bj[b];
          acc[a][b] = fmaf(df, df, acc[a][b]);
        }
    }
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------- pass 1: row sums
__global__ void __launch_bounds__(kNdivThreads)
ndiv_rowsum_kernel(const float* __restrict__ x, const float* __restrict__ z, int k, int Dx, int Dz,
                   int row0, int rows, float* __restrict__ sx, float* __restrict__ sz) {
  __shared__ __align__(16) float s_i[kDC * kPadI];
  __shared__ __align__(16) float s_j[kDC * kPadJ];
  const int n = blockIdx.y;
  const int i0 = row0 + blockIdx.x * kTI;
  const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  float rsx[4] = {0, 0, 0, 0}, rsz[4] = {0, 0, 0, 0};
  float acc[4][4];
  for (int j0 = 0; j0 < k; j0 += kTJ) {
    sqdist_tile(acc, xg, k, Dx, i0, j0, s_i, s_j);
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b)
        if (j0 + tx * 4 + b < k) rsx[a] += sqrtf(acc[a][b]);
    sqdist_tile(acc, zg, k, Dz, i0, j0, s_i, s_j);
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b)
        if (j0 + tx * 4 + b < k) rsz[a] += sqrtf(acc[a][b]);
  }
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const float vx = warp_sum(rsx[a]), vz = warp_sum(rsz[a]);
    const int i = i0 + ty * 4 + a;
    if (tx == 0 && i < k && i < row0 + rows) {
      sx[(size_t)n * k + i] = vx;
      sz[(size_t)n * k + i] = vz;
    }
  }
}

// ---------------------------------------------------------------- pass 2: loss + gradient
// NCH = number of 16-wide feature chunks of x handled per dimension pass (grad accumulators)
template <int NCH, bool WITH_GRAD>
__global__ void __launch_bounds__(kNdivThreads)
ndiv_loss_grad_kernel(const float* __restrict__ x, const float* __restrict__ z,
                      const float* __restrict__ sx, const float* __restrict__ sz, int k, int Dx, int Dz,
                      float margin, int row0, int rows, int dpass0, float* __restrict__ loss_part,
                      float* __restrict__ grad_x) {
  __shared__ __align__(16) float s_i[kDC * kPadI];
  __shared__ __align__(16) float s_j[kDC * kPadJ];
  __shared__ __align__(16) float s_W[kTI * kPadJ];
  __shared__ float s_red[kNdivThreads / 32];
  const int n = blockIdx.y;
  const int i0 = row0 + blockIdx.x * kTI;
  const int iend = (row0 + rows) < k ? (row0 + rows) : k;
  const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
  const int gi = threadIdx.x >> 3, dg = threadIdx.x & 7;
  const float* xg = x + (size_t)n * k * Dx;
  const float* zg = z + (size_t)n * k * Dz;
  const float* sxg = sx + (size_t)n * k;
  const float* szg = sz + (size_t)n * k;

  float sxi[4], szi[4];
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int i = i0 + ty * 4 + a;
    sxi[a] = i < k ? __ldg(sxg + i) : 1.f;
    szi[a] = i < k ? __ldg(szg + i) : 1.f;
  }
  float gacc[NCH * 2];
#pragma unroll
  for (int q = 0; q < NCH * 2; ++q) gacc[q] = 0.f;
  float sumW = 0.f, lsum = 0.f;
  float dx[4][4], dz[4][4];

  for (int j0 = 0; j0 < k; j0 += kTJ) {
    sqdist_tile(dx, xg, k, Dx, i0, j0, s_i, s_j);
    sqdist_tile(dz, zg, k, Dz, i0, j0, s_i, s_j);
    float sxj[4], szj[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int j = j0 + tx * 4 + b;
      sxj[b] = j < k ? __ldg(sxg + j) : 1.f;
      szj[b] = j < k ? __ldg(szg + j) : 1.f;
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int i = i0 + ty * 4 + a;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int j = j0 + tx * 4 + b;
        const float ddx = sqrtf(dx[a][b]), ddz = sqrtf(dz[a][b]);
        // (i,j): relu(margin * dz/sz_i - dx/sx_i), diversity.py:18,40
        const float hij = __fsub_rn(__fmul_rn(margin, ddz / szi[a]), ddx / sxi[a]);
        const float hji = __fsub_rn(__fmul_rn(margin, ddz / szj[b]), ddx / sxj[b]);
        float w = 0.f;
        if (i < iend && j < k) {
          if (dpass0 == 0) lsum += (hij > 0.f) ? hij : ((hij != hij) ? hij : 0.f);
          const float wij = (hij > 0.f ? 1.f / sxi[a] : 0.f) + (hji > 0.f ? 1.f / sxj[b] : 0.f);
          w = ddx > 0.f ? -wij / ddx : 0.f;
        }
        if (WITH_GRAD) s_W[(ty * 4 + a) * kPadJ + tx * 4 + b] = w;
      }
    }
    if (WITH_GRAD) {
      // grad_i -= sum_j What_ij x_j  (a [32 x 128] x [128 x Dx] product), sumW_i = sum_j What_ij
#pragma unroll
      for (int cc = 0; cc < NCH; ++cc) {
        const int d0 = dpass0 + cc * kDC;
        if (d0 < Dx) {
          __syncthreads();
          stage_chunk<kTJ, kPadJ>(s_j, xg, k, Dx, j0, d0);
          __syncthreads();
          const float* wrow = s_W + gi * kPadJ;
          const float* xa = s_j + (dg * 2) * kPadJ;
          const float* xb = xa + kPadJ;
          float g0 = 0.f, g1 = 0.f, sw = 0.f;
#pragma unroll 8
          for (int j = 0; j < kTJ; j += 4) {
            const float4 w4 = *reinterpret_cast<const float4*>(wrow + j);
            const float4 a4 = *reinterpret_cast<const float4*>(xa + j);
            const float4 b4 = *reinterpret_cast<const float4*>(xb + j);
            g0 = fmaf(w4.x, a4.x, g0); g0 = fmaf(w4.y, a4.y, g0); g0 = fmaf(w4.z, a4.z, g0); g0 = fmaf(w4.w, a4.w, g0);
            g1 = fmaf(w4.x, b4.x, g1); g1 = fmaf(w4.y, b4.y, g1); g1 = fmaf(w4.z, b4.z, g1); g1 = fmaf(w4.w, b4.w, g1);
            sw += (w4.x + w4.y) + (w4.z + w4.w);
          }
          gacc[cc * 2] += g0;
          gacc[cc * 2 + 1] += g1;
          if (cc == 0) sumW += sw;
        }
      }
    }
  }
  if (WITH_GRAD) {
    const int i = i0 + gi;
    if (i < iend) {
#pragma unroll
      for (int cc = 0; cc < NCH; ++cc) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int d = dpass0 + cc * kDC + dg * 2 + h;
          if (d < Dx) {
            const size_t off = ((size_t)n * k + i) * Dx + d;
            grad_x[off] = __ldg(x + off) * sumW - gacc[cc * 2 + h];
          }
        }
      }
    }
  }
  if (dpass0 == 0) {
    lsum = warp_sum(lsum);
    __syncthreads();
    if (tx == 0) s_red[ty] = lsum;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kNdivThreads / 32; ++w) t += s_red[w];
      loss_part[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
  }
}

__global__ void ndiv_finalize_kernel(const float* __restrict__ part, int n, float* __restrict__ loss) {
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

extern "C" int ndp_ndiv_workspace_bytes(int N, int k, size_t* bytes) {
  NDP_CHECK_ARG(bytes && N >= 0 && k >= 0, "ndiv: bad N/k");
  *bytes = (size_t)N * ((k + kTI - 1) / kTI + 1) * sizeof(float) + 256;
  return NDP_OK;
}

static int ndiv_args_ok(const float* x, const float* z, int N, int k, int Dx, int Dz, int row0, int rows) {
  NDP_CHECK_ARG(N >= 0 && k >= 0 && Dx >= 1 && Dz >= 1, "ndiv: bad shape N=%d k=%d Dx=%d Dz=%d", N, k, Dx, Dz);
  NDP_CHECK_ARG(N <= 65535, "ndiv: N must be <= 65535 per call");
  NDP_CHECK_ARG(row0 >= 0 && rows >= 0 && row0 + rows <= k, "ndiv: bad row block [%d,+%d) of %d", row0, rows, k);
  NDP_CHECK_ARG(N == 0 || k == 0 || (x && z), "ndiv: null x/z");
  return NDP_OK;
}

extern "C" int ndp_ndiv_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz,
                                int row0, int rows, float* sx, float* sz, void* stream) {
  if (int rc = ndiv_args_ok(x, z, N, k, Dx, Dz, row0, rows)) return rc;
  if (int rc = device_check()) return rc;
  if (N == 0 || rows == 0) return NDP_OK;
  NDP_CHECK_ARG(sx && sz, "ndiv: null sx/sz");
  dim3 grid((rows + kTI - 1) / kTI, N);
  ndiv_rowsum_kernel<<<grid, kNdivThreads, 0, (cudaStream_t)stream>>>(x, z, k, Dx, Dz, row0, rows, sx, sz);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

template <int NCH>
static void launch_loss_grad(dim3 grid, cudaStream_t s, const float* x, const float* z, const float* sx,
                             const float* sz, int k, int Dx, int Dz, float margin, int row0, int rows,
                             int dpass0, float* part, float* grad_x) {
  if (grad_x)
    ndiv_loss_grad_kernel<NCH, true><<<grid, kNdivThreads, 0, s>>>(x, z, sx, sz, k, Dx, Dz, margin, row0, rows,
                                                                  dpass0, part, grad_x);
  else
    ndiv_loss_grad_kernel<NCH, false><<<grid, kNdivThreads, 0, s>>>(x, z, sx, sz, k, Dx, Dz, margin, row0, rows,
                                                                   dpass0, part, grad_x);
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
  dim3 grid((rows + kTI - 1) / kTI, N);
  const size_t nparts = (size_t)grid.x * grid.y;
  if (!ws || ws_bytes < nparts * sizeof(float)) {
    set_error("ndiv: workspace %zu < %zu", ws_bytes, nparts * sizeof(float));
    return NDP_ERR_WORKSPACE;
  }
  float* part = reinterpret_cast<float*>(ws);
  const int per_pass = 16 * kDC;
  for (int dp = 0; dp < Dx; dp += per_pass) {
    const int chunks = ((Dx - dp < per_pass ? Dx - dp : per_pass) + kDC - 1) / kDC;
    if (!grad_x || chunks <= 1) launch_loss_grad<1>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, dp, part, grad_x);
    else if (chunks <= 2) launch_loss_grad<2>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, dp, part, grad_x);
    else if (chunks <= 4) launch_loss_grad<4>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, dp, part, grad_x);
    else if (chunks <= 8) launch_loss_grad<8>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, dp, part, grad_x);
    else launch_loss_grad<16>(grid, s, x, z, sx, sz, k, Dx, Dz, margin, row0, rows, dp, part, grad_x);
    NDP_LAUNCH_CHECK();
    if (!grad_x) break;
  }
  ndiv_finalize_kernel<<<1, 256, 0, s>>>(part, (int)nparts, loss);
  NDP_LAUNCH_CHECK();
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
    ndiv_rowsum_kernel<<<dim3((k + kTI - 1) / kTI, N), kNdivThreads, 0, (cudaStream_t)stream>>>(
        v, v, k, C, C, 0, k, rowsum_ws, rowsum_ws);
    NDP_LAUNCH_CHECK();
  }
  pairwise_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(v, grad_out, k, C, normalise,
                                                             normalise ? rowsum_ws : v, grad_v);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
