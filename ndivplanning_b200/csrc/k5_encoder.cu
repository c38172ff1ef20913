// K5 — image-encoder front end (first slice): models/image_autoencoder.py:14-49 in eval mode,
//   five 3x3 stride-2 pad-1 convolutions 3 -> 64 -> 128 -> 256 -> 512 -> 1024 (BatchNorm on the first three, ReLU
//   after each), then a 4x4 valid convolution 1024 -> 128:   [B,3,128,128] -> [B,128,1,1].
// It is what turns camera frames into the embeddings the planner rolls (mpc_eval.py:135-136) and the GAN step
// conditions on (train_gan.py:152-155); the reference re-encodes the goal K*T times per plan — here once.
//
// Every layer is an implicit GEMM  Y[m, n] = sum_k A(m, k) W[n, k],  m = (image, out row, out col),
// k = (kernel row, kernel col, in channel): the im2col matrix is never materialised, the loader gathers it (zero
// padding = zero-fill async copies).  Activations are NHWC between layers (channels contiguous = contiguous k);
// the first layer reads the reference's NCHW images directly.  Eval-mode BatchNorm is folded into the weights and
// bias when the parameters are loaded.  Small batches (a plan encodes two frames) leave few output tiles, so the
// contraction is split over blockIdx.z into partial slots that a finalize kernel adds in a fixed order (+ bias,
// ReLU): deterministic, no atomics.  Exact fp32 FFMA in this slice; the tcgen05 variant follows the K4 kernels.
#include <vector>
#include "common.cuh"
#include "encoder_tc.cuh"

namespace ndp {
namespace enc {

constexpr int kTM = 64, kTN = 64, kTK = 16, kThreads = 256, kStages = 4;

struct ConvP {
  const float* in;      // NCHW (layer 1) or NHWC
  int nchw;
  int B, H, W, Cin;     // input geometry
  int OH, OW, Cout;
  int KH, KW, stride, pad;
  const float* Wt;      // [Cout][KH*KW*Cin], k = (kh*KW + kw)*Cin + ci
  float* part;          // [splits][M][Cout]
  int M, K, k_per_split;
  float* amax;          // nullable [1]: running maximum of the direct (fused-epilogue) outputs, for the tcgen05 layers
};

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(kThreads) conv_gemm_kernel(const ConvP p, const float* __restrict__ bias, int relu,
                                                             float* __restrict__ direct_out) {
  __shared__ __align__(16) float As[kStages][kTK][kTM + 4];
  __shared__ __align__(16) float Bs[kStages][kTK][kTN + 4];
  const int t = threadIdx.x;
  const int m0 = blockIdx.x * kTM, n0 = blockIdx.y * kTN;
  const int k_begin = blockIdx.z * p.k_per_split;
  const int k_end = min(p.K, k_begin + p.k_per_split);
  const int ti = t >> 4, tj = t & 15;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0.f;

  // this thread stages rows i = (t >> 4) + 16 q of the A tile: their (image, out row, out col) never change
  int ib[4], ih0[4], iw0[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int m = m0 + (t >> 4) + 16 * q;
    if (m < p.M) {
      const int b = m / (p.OH * p.OW), r = m % (p.OH * p.OW);
      ib[q] = b; ih0[q] = (r / p.OW) * p.stride - p.pad; iw0[q] = (r % p.OW) * p.stride - p.pad;
    } else {
      ib[q] = -1; ih0[q] = iw0[q] = 0;
    }
  }
  auto issue = [&](int stage, int k0) {
    const int k = k0 + (t & 15);
    int ci = 0, kh = 0, kw = 0;
    const bool kok = k < k_end;
    if (kok) { ci = k % p.Cin; const int r = k / p.Cin; kw = r % p.KW; kh = r / p.KW; }
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int ih = ih0[q] + kh, iw = iw0[q] + kw;
      const bool ok = kok && ib[q] >= 0 && ih >= 0 && ih < p.H && iw >= 0 && iw < p.W;
      const float* src = p.in;
      if (ok) src += p.nchw ? (((size_t)ib[q] * p.Cin + ci) * p.H + ih) * p.W + iw
                            : (((size_t)ib[q] * p.H + ih) * p.W + iw) * p.Cin + ci;
      cp_async4(&As[stage][t & 15][(t >> 4) + 16 * q], src, ok);
    }
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int n = n0 + (t >> 4) + 16 * q;
      const bool ok = kok && n < p.Cout;
      cp_async4(&Bs[stage][t & 15][(t >> 4) + 16 * q], ok ? p.Wt + (size_t)n * p.K + k : p.Wt, ok);
    }
  };

  const int nchunks = (k_end - k_begin + kTK - 1) / kTK;
#pragma unroll
  for (int st = 0; st < kStages - 1; st++) {
    if (st < nchunks) issue(st, k_begin + st * kTK);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    const int st = c % kStages;
    cp_async_wait<kStages - 2>();
    __syncthreads();
    if (c + kStages - 1 < nchunks) issue((c + kStages - 1) % kStages, k_begin + (c + kStages - 1) * kTK);
    cp_async_commit();
#pragma unroll
    for (int kk = 0; kk < kTK; kk++) {
      const float4 av = *reinterpret_cast<const float4*>(&As[st][kk][ti * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[st][kk][tj * 4]);
      const float a[4] = {av.x, av.y, av.z, av.w}, b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int x = 0; x < 4; x++)
#pragma unroll
        for (int y = 0; y < 4; y++) acc[x][y] = fmaf(a[x], b[y], acc[x][y]);
    }
  }
  float* out = direct_out ? direct_out : p.part + (size_t)blockIdx.z * p.M * p.Cout;   // direct: one split, epilogue here
  float mx = 0.f;
#pragma unroll
  for (int x = 0; x < 4; x++) {
    const int m = m0 + ti * 4 + x;
    if (m >= p.M) continue;
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const int n = n0 + tj * 4 + y;
      if (n >= p.Cout) continue;
      float v = acc[x][y];
      if (direct_out) { v += __ldg(bias + n); if (relu) v = fmaxf(v, 0.f); mx = fmaxf(mx, v); }
      out[(size_t)m * p.Cout + n] = v;
    }
  }
  if (direct_out && p.amax) {                              // post-ReLU values: non-negative, ordered like their bits
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((t & 31) == 0 && mx > 0.f) atomicMax(reinterpret_cast<int*>(p.amax), __float_as_int(mx));
  }
}

// Large-batch variant for the NHWC layers (Cin % 16 == 0: a 16-wide k chunk stays inside one kernel cell, so its 16
// input channels are contiguous and (kh, kw) is uniform over the CTA).  128 x 128 tile, 8 x 8 outputs per thread
// (rows ti + 16 x, columns tj + 16 y: conflict-free 16-B shared loads with a 20-float row pitch), 16-B zero-fill
// async copies, 4 stages, 2 CTAs per SM.  With one split the bias + ReLU epilogue writes the activation directly.
constexpr int kBM = 128, kBN = 128, kBLD = 20, kBStages = 4;
constexpr int kBStageFloats = (kBM + kBN) * kBLD;
constexpr int kBSmem = kBStages * kBStageFloats * 4;

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}

__global__ void __launch_bounds__(kThreads, 2)
conv_gemm_big_kernel(const ConvP p, const float* __restrict__ bias, int relu, float* __restrict__ direct_out) {
  extern __shared__ __align__(16) float smem[];
  const int t = threadIdx.x;
  const int m0 = blockIdx.x * kBM, n0 = blockIdx.y * kBN;
  const int k_begin = blockIdx.z * p.k_per_split;
  const int k_end = min(p.K, k_begin + p.k_per_split);
  const int ti = t >> 4, tj = t & 15;
  const int lr = t >> 2, lv = (t & 3) * 4;          // loader: rows lr, lr + 64; floats lv..lv+3 of the chunk
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; a++)
#pragma unroll
    for (int b = 0; b < 8; b++) acc[a][b] = 0.f;

  int ih0[2], iw0[2];
  const float* abase[2];
  const float* wbase[2];
  bool wok[2];
#pragma unroll
  for (int q = 0; q < 2; q++) {
    const int m = m0 + lr + 64 * q;
    if (m < p.M) {
      const int b = m / (p.OH * p.OW), r = m % (p.OH * p.OW);
      ih0[q] = (r / p.OW) * p.stride - p.pad; iw0[q] = (r % p.OW) * p.stride - p.pad;
      abase[q] = p.in + (size_t)b * p.H * p.W * p.Cin + lv;
    } else {
      ih0[q] = -(1 << 20); iw0[q] = 0; abase[q] = p.in;
    }
    const int n = n0 + lr + 64 * q;
    wok[q] = n < p.Cout;
    wbase[q] = wok[q] ? p.Wt + (size_t)n * p.K + lv : p.Wt;
  }
  auto issue = [&](int stage, int k0) {
    float* As = smem + stage * kBStageFloats;
    float* Bs = As + kBM * kBLD;
    const int cell = k0 / p.Cin, ci0 = k0 - cell * p.Cin;
    const int kh = cell / p.KW, kw = cell - kh * p.KW;
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const int ih = ih0[q] + kh, iw = iw0[q] + kw;
      const bool ok = ih >= 0 && ih < p.H && iw >= 0 && iw < p.W;
      cp_async16(As + (lr + 64 * q) * kBLD + lv, ok ? abase[q] + ((size_t)ih * p.W + iw) * p.Cin + ci0 : p.in, ok);
      cp_async16(Bs + (lr + 64 * q) * kBLD + lv, wok[q] ? wbase[q] + k0 : p.Wt, wok[q]);
    }
  };

  const int nchunks = (k_end - k_begin) / kTK;
#pragma unroll
  for (int st = 0; st < kBStages - 1; st++) {
    if (st < nchunks) issue(st, k_begin + st * kTK);
    cp_async_commit();
  }
  for (int c = 0; c < nchunks; c++) {
    cp_async_wait<kBStages - 2>();
    __syncthreads();
    if (c + kBStages - 1 < nchunks) issue((c + kBStages - 1) % kBStages, k_begin + (c + kBStages - 1) * kTK);
    cp_async_commit();
    const float* As = smem + (c % kBStages) * kBStageFloats;
    const float* Bs = As + kBM * kBLD;
#pragma unroll 1
    for (int kq = 0; kq < 4; kq++) {
      float4 a[8];
#pragma unroll
      for (int x = 0; x < 8; x++) a[x] = *reinterpret_cast<const float4*>(As + (ti + 16 * x) * kBLD + kq * 4);
#pragma unroll
      for (int y = 0; y < 8; y++) {
        const float4 b = *reinterpret_cast<const float4*>(Bs + (tj + 16 * y) * kBLD + kq * 4);
#pragma unroll
        for (int x = 0; x < 8; x++) {
          float v = acc[x][y];
          v = fmaf(a[x].x, b.x, v); v = fmaf(a[x].y, b.y, v); v = fmaf(a[x].z, b.z, v); v = fmaf(a[x].w, b.w, v);
          acc[x][y] = v;
        }
      }
    }
  }
  float* out = direct_out ? direct_out : p.part + (size_t)blockIdx.z * p.M * p.Cout;
#pragma unroll
  for (int x = 0; x < 8; x++) {
    const int m = m0 + ti + 16 * x;
    if (m >= p.M) continue;
#pragma unroll
    for (int y = 0; y < 8; y++) {
      const int n = n0 + tj + 16 * y;
      if (n >= p.Cout) continue;
      float v = acc[x][y];
      if (direct_out) { v += __ldg(bias + n); if (relu) v = fmaxf(v, 0.f); }
      out[(size_t)m * p.Cout + n] = v;
    }
  }
}

// out[m][n] = act( sum_s part[s][m][n] + bias[n] ), splits added in order
__global__ void __launch_bounds__(256) conv_finalize_kernel(const float* __restrict__ part, int splits, size_t MN, int N,
                                                            const float* __restrict__ bias, int relu,
                                                            float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= MN) return;
  float v = 0.f;
  for (int s = 0; s < splits; s++) v += part[(size_t)s * MN + i];
  v += __ldg(bias + i % N);
  out[i] = relu ? fmaxf(v, 0.f) : v;
}

// same, for many splits over few outputs (conv6: up to 147 splits): a block takes 32 consecutive outputs, warp w adds
// splits w, w + 8, ... (coalesced 128-B rows), then the 8 warp sums are added in warp order
__global__ void __launch_bounds__(256) conv_finalize_wide_kernel(const float* __restrict__ part, int splits, size_t MN,
                                                                 int N, const float* __restrict__ bias, int relu,
                                                                 float* __restrict__ out) {
  __shared__ float sm[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t i = (size_t)blockIdx.x * 32 + lane;
  float v = 0.f;
  if (i < MN)
    for (int s = w; s < splits; s += 8) v += part[(size_t)s * MN + i];
  sm[w][lane] = v;
  __syncthreads();
  if (w == 0 && i < MN) {
    float t = sm[0][lane];
#pragma unroll
    for (int k = 1; k < 8; k++) t += sm[k][lane];
    t += __ldg(bias + i % N);
    out[i] = relu ? fmaxf(t, 0.f) : t;
  }
}

struct Layer {
  int Cin, Cout, KH, KW, stride, pad, H, W, OH, OW, relu;
  size_t w_off, b_off;
};

}  // namespace enc
}  // namespace ndp

// ---- handle ------------------------------------------------------------------------------------------------
using namespace ndp;
using namespace ndp::enc;

struct ndp_encoder {
  int max_B, L;
  Layer ly[6];
  void* arena;
  float* params;        // folded weights [Cout][KH][KW][Cin] and biases, all layers
  float* act[2];        // NHWC activations, ping-pong
  float* part;          // split-K partial slots
  size_t part_floats;
  // tcgen05 path of conv2..conv5 (k5_encoder_tc.cu), batches of >= kTcMinB images
  __half* act_hi;       // scaled fp16 planes of the current layer's input
  __half* act_lo;
  uint8_t* wpack[6];    // MMA-ready weight tiles (layers 1..4)
  float* wsinv[6];      // per-output-channel 1 / weight scale
  float* ws_tmp;        // [1024] weight scales during packing
  float* amax;          // [8] per-layer running maximum of the layer's INPUT activation
  float* ainv;          // [8] per-layer 1 / activation scale
};
static constexpr int kTcMinB = 16;

static int plan_splits(int M, int Cout, int K, size_t part_floats) {
  const int tiles = ((M + kTM - 1) / kTM) * ((Cout + kTN - 1) / kTN);
  int s = (2 * num_sms() + tiles - 1) / tiles;
  const int max_by_k = (K + 4 * kTK - 1) / (4 * kTK);       // at least 4 chunks per split
  if (s > max_by_k) s = max_by_k;
  while (s > 1 && (size_t)s * M * Cout > part_floats) --s;
  return s < 1 ? 1 : s;
}

extern "C" {

// geometry of models/image_autoencoder.py:17-27 for img x img inputs (128 in the reference)
int ndp_encoder_create(ndp_encoder** out, int max_images, int img) {
  NDP_CHECK_ARG(out && max_images > 0 && img >= 128 && img % 32 == 0 && img / 32 == 4,
                "ndp_encoder_create: the 4x4 valid conv6 needs img / 32 == 4 (img = 128)");
  if (int rc = device_check()) return rc;
  ndp_encoder* e = new ndp_encoder();
  e->max_B = max_images; e->L = 6;
  const int ch[6] = {3, 64, 128, 256, 512, 1024};
  size_t off = 0, max_act = 0, max_mn = 0;
  int h = img;
  for (int l = 0; l < 6; l++) {
    Layer& y = e->ly[l];
    if (l < 5) { y.Cin = ch[l]; y.Cout = ch[l + 1]; y.KH = y.KW = 3; y.stride = 2; y.pad = 1; y.relu = 1; }
    else { y.Cin = 1024; y.Cout = 128; y.KH = y.KW = 4; y.stride = 1; y.pad = 0; y.relu = 0; }
    y.H = y.W = h;
    y.OH = y.OW = (h + 2 * y.pad - y.KH) / y.stride + 1;
    h = y.OH;
    y.w_off = off; off += (size_t)y.Cout * y.KH * y.KW * y.Cin;
    y.b_off = off; off += (size_t)y.Cout;
    off = (off + 3) & ~(size_t)3;
    const size_t mn = (size_t)max_images * y.OH * y.OW * y.Cout;
    max_act = mn > max_act ? mn : max_act;
    max_mn = mn > max_mn ? mn : max_mn;
  }
  e->part_floats = max_mn * 4 > ((size_t)16 << 20) ? max_mn * 4 : ((size_t)16 << 20);   // >= 64 MB of partial slots
  size_t tc_floats = 0;                                    // fp16 planes (2 x max_act halves) + packed weights + scales
  if (max_images >= kTcMinB) {
    tc_floats = max_act + 2048 + 64;
    for (int l = 1; l <= 4; l++) tc_floats += conv_tc_pack_bytes(e->ly[l].Cout, e->ly[l].KH * e->ly[l].KW * e->ly[l].Cin) / 4 + e->ly[l].Cout + 64;
  }
  const size_t bytes = (off + 2 * max_act + e->part_floats + tc_floats) * sizeof(float) + 4096;
  cudaError_t err = cudaMalloc(&e->arena, bytes);
  if (err != cudaSuccess) {
    set_error("ndp_encoder_create: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(err));
    delete e;
    return NDP_ERR_CUDA;
  }
  e->params = (float*)e->arena;
  e->act[0] = e->params + ((off + 63) & ~(size_t)63);
  e->act[1] = e->act[0] + max_act;
  e->part = e->act[1] + max_act;
  e->part_floats -= 64;
  e->act_hi = nullptr;
  if (tc_floats) {
    float* q = e->part + e->part_floats + 64;
    q = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(q) + 255) & ~(uintptr_t)255);
    e->act_hi = reinterpret_cast<__half*>(q);
    e->act_lo = e->act_hi + max_act;
    q += max_act;
    e->ws_tmp = q; q += 1024;
    e->amax = q; e->ainv = q + 8; q += 64;
    for (int l = 1; l <= 4; l++) {
      const Layer& y = e->ly[l];
      q = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(q) + 255) & ~(uintptr_t)255);
      e->wpack[l] = reinterpret_cast<uint8_t*>(q);
      q += conv_tc_pack_bytes(y.Cout, y.KH * y.KW * y.Cin) / 4;
      e->wsinv[l] = q; q += y.Cout;
    }
  }
  *out = e;
  return NDP_OK;
}

void ndp_encoder_destroy(ndp_encoder* e) {
  if (!e) return;
  cudaFree(e->arena);
  delete e;
}

// W[l]: device pointer to layer l's weights ALREADY in [Cout][KH][KW][Cin] order with eval-mode BatchNorm folded in
// (W' = W * gamma / sqrt(var + eps), b' = (b - mean) * gamma / sqrt(var + eps) + beta); b[l]: folded bias [Cout].
int ndp_encoder_load(ndp_encoder* e, const float* const* W, const float* const* b, void* stream) {
  NDP_CHECK_ARG(e && W && b, "ndp_encoder_load: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  for (int l = 0; l < e->L; l++) {
    const Layer& y = e->ly[l];
    NDP_CUDA(cudaMemcpyAsync(e->params + y.w_off, W[l], (size_t)y.Cout * y.KH * y.KW * y.Cin * 4, cudaMemcpyDeviceToDevice, s));
    NDP_CUDA(cudaMemcpyAsync(e->params + y.b_off, b[l], (size_t)y.Cout * 4, cudaMemcpyDeviceToDevice, s));
  }
  if (e->act_hi)
    for (int l = 1; l <= 4; l++) {
      const Layer& y = e->ly[l];
      if (int rc = conv_tc_pack(e->params + y.w_off, y.Cout, y.KH * y.KW * y.Cin, e->wpack[l], e->wsinv[l], e->ws_tmp, s)) return rc;
    }
  return NDP_OK;
}

// images [B,3,img,img] NCHW fp32 (device) -> codes [B,128] (== the reference's [B,128,1,1])
int ndp_encoder_forward(ndp_encoder* e, const float* images, int B, float* codes, void* stream) {
  NDP_CHECK_ARG(e && images && codes, "ndp_encoder_forward: null argument");
  NDP_CHECK_ARG(B > 0 && B <= e->max_B, "ndp_encoder_forward: B=%d outside (0, %d]", B, e->max_B);
  cudaStream_t s = (cudaStream_t)stream;
  const float* in = images;
  static const bool tc_off = getenv("NDP_ENC_NO_TC") != nullptr;      // diagnostics: force the CUDA-core kernels
  const bool use_tc = e->act_hi && !tc_off && B >= kTcMinB;
  if (use_tc) NDP_CUDA(cudaMemsetAsync(e->amax, 0, 8 * sizeof(float), s));
  for (int l = 0; l < e->L; l++) {
    const Layer& y = e->ly[l];
    if (use_tc && l >= 1 && l <= 4) {
      // tcgen05 layer: split the fp32 input into scaled fp16 hi / lo planes, implicit GEMM, fp32 NHWC output
      ConvTcP t{};
      t.B = B; t.H = y.H; t.W = y.W; t.Cin = y.Cin; t.OH = y.OH; t.OW = y.OW; t.Cout = y.Cout; t.KW = y.KW;
      t.stride = y.stride; t.pad = y.pad; t.M = B * y.OH * y.OW; t.K = y.KH * y.KW * y.Cin;
      if (int rc = conv_tc_split(in, (size_t)B * y.H * y.W * y.Cin, e->amax + l, e->ainv + l, e->act_hi, e->act_lo, s)) return rc;
      t.in_hi = e->act_hi; t.in_lo = e->act_lo; t.wpack = e->wpack[l]; t.wsinv = e->wsinv[l]; t.ainv = e->ainv + l;
      t.bias = e->params + y.b_off; t.relu = y.relu;
      t.out = e->act[l & 1]; t.part = e->part; t.amax_out = l < 4 ? e->amax + l + 1 : nullptr;
      if (int rc = conv_tc_forward(t, e->part_floats, s)) return rc;
      in = t.out;
      continue;
    }
    ConvP p;
    p.amax = (use_tc && l == 0) ? e->amax + 1 : nullptr;
    p.in = in; p.nchw = l == 0;
    p.B = B; p.H = y.H; p.W = y.W; p.Cin = y.Cin; p.OH = y.OH; p.OW = y.OW; p.Cout = y.Cout;
    p.KH = y.KH; p.KW = y.KW; p.stride = y.stride; p.pad = y.pad;
    p.Wt = e->params + y.w_off;
    p.part = e->part;
    p.M = B * y.OH * y.OW; p.K = y.KH * y.KW * y.Cin;
    float* out = l == e->L - 1 ? codes : e->act[l & 1];
    const size_t MN = (size_t)p.M * p.Cout;
    const int big_tiles = ((p.M + kBM - 1) / kBM) * ((p.Cout + kBN - 1) / kBN);
    if (!p.nchw && p.Cin % kTK == 0 && big_tiles >= 32) {
      static std::atomic<unsigned long long> attr_devs{0};
      if (first_use_on_device(attr_devs)) {
        NDP_CUDA(cudaFuncSetAttribute(conv_gemm_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBSmem));
      }
      int splits = (2 * num_sms() + big_tiles - 1) / big_tiles;      // two resident CTAs per SM
      if (big_tiles >= num_sms()) splits = 1;
      const int max_by_k = p.K / (8 * kTK) > 0 ? p.K / (8 * kTK) : 1;
      if (splits > max_by_k) splits = max_by_k;
      while (splits > 1 && (size_t)splits * MN > e->part_floats) --splits;
      p.k_per_split = ((p.K + splits - 1) / splits + kTK - 1) / kTK * kTK;
      const int zs = (p.K + p.k_per_split - 1) / p.k_per_split;
      conv_gemm_big_kernel<<<dim3((p.M + kBM - 1) / kBM, (p.Cout + kBN - 1) / kBN, zs), kThreads, kBSmem, s>>>(
          p, e->params + y.b_off, y.relu, zs == 1 ? out : nullptr);
      NDP_LAUNCH_CHECK();
      if (zs > 1) {
        conv_finalize_kernel<<<(unsigned)((MN + 255) / 256), 256, 0, s>>>(e->part, zs, MN, p.Cout,
                                                                          e->params + y.b_off, y.relu, out);
        NDP_LAUNCH_CHECK();
      }
      in = out;
      continue;
    }
    const int splits = plan_splits(p.M, p.Cout, p.K, e->part_floats);
    p.k_per_split = ((p.K + splits - 1) / splits + kTK - 1) / kTK * kTK;
    const int zs = (p.K + p.k_per_split - 1) / p.k_per_split;
    conv_gemm_kernel<<<dim3((p.M + kTM - 1) / kTM, (p.Cout + kTN - 1) / kTN, zs), kThreads, 0, s>>>(
        p, e->params + y.b_off, y.relu, zs == 1 ? out : nullptr);
    NDP_LAUNCH_CHECK();
    if (zs == 1) { in = out; continue; }
    if (zs >= 16)
      conv_finalize_wide_kernel<<<(unsigned)((MN + 31) / 32), 256, 0, s>>>(e->part, zs, MN, p.Cout,
                                                                          e->params + y.b_off, y.relu, out);
    else
      conv_finalize_kernel<<<(unsigned)((MN + 255) / 256), 256, 0, s>>>(e->part, zs, MN, p.Cout, e->params + y.b_off,
                                                                        y.relu, out);
    NDP_LAUNCH_CHECK();
    in = out;
  }
  return NDP_OK;
}

}  // extern "C"
