// C-ABI glue of libndivplan: error reporting, device gate, planner handle (weight prepack),
// the single-call generator / forward-model entry points, launch counting.
#include <stdarg.h>
#include <string.h>
#include "planner.cuh"

namespace ndp {

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int device_check() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_rc = 0;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    set_error("no CUDA device: libndivplan has no CPU path");
    return NDP_ERR_DEVICE;
  }
  if (dev == cached_dev) return cached_rc;
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  cached_dev = dev;
  cached_rc = (major == 10) ? NDP_OK : NDP_ERR_DEVICE;
  if (cached_rc) set_error("device %d is sm_%d%d; libndivplan is built for sm_100a only", dev, major, minor);
  return cached_rc;
}

int num_sms() {
  int dev = 0, n = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  return n;
}

// dst[kk*N + n] = src[(n0+n)*ld + k0 + kk]   (transpose-pack a [N x Kc] block of a torch [out,in] matrix)
__global__ void pack_T_kernel(float* __restrict__ dst, const float* __restrict__ src, int ld, int n0, int k0,
                              int Kc, int N) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Kc * N) return;
  const int kk = idx / N, n = idx % N;
  dst[idx] = src[(size_t)(n0 + n) * ld + k0 + kk];
}

// b1g[g][n] = b1[n] + sum_e W1g[n][e] * goal[g][e]   — goal half of generator fc1, once per plan
// (the reference recomputes it for every candidate and step: mpc_eval.py:133,136-138)
__global__ void goal_bias_kernel(const float* __restrict__ w1g, const float* __restrict__ b1,
                                 const float* __restrict__ goal, int E, float* __restrict__ b1g) {
  const int g = blockIdx.x;
  const int n = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= 128) return;
  // sequential-order friendly: each lane sums a strided subset, then tree
  float acc = 0.f;
  for (int e = lane; e < E; e += 32) acc = fmaf(__ldg(w1g + (size_t)n * E + e), __ldg(goal + (size_t)g * E + e), acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) b1g[(size_t)g * 128 + n] = acc + __ldg(b1 + n);
}

// out[r][n] = act( b[n] + sum_k in[r][k] W[n][k] ) (+ res[r][n]); one warp per output.
// in rows are formed from up to two sources (concat): in0 [B,K0] then in1 [B,K1].
__global__ void linear_kernel(const float* __restrict__ in0, int K0, const float* __restrict__ in1, int K1,
                              const float* __restrict__ W, const float* __restrict__ b, int N, int B, int relu,
                              const float* __restrict__ res, float* __restrict__ out) {
  const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (wid >= (long long)B * N) return;
  const int r = (int)(wid / N), n = (int)(wid % N);
  const int K = K0 + K1;
  const float* w = W + (size_t)n * K;
  float acc = 0.f;
  for (int k = lane; k < K0; k += 32) acc = fmaf(__ldg(in0 + (size_t)r * K0 + k), __ldg(w + k), acc);
  for (int k = lane; k < K1; k += 32) acc = fmaf(__ldg(in1 + (size_t)r * K1 + k), __ldg(w + K0 + k), acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) {
    float v = acc + __ldg(b + n);
    if (relu) v = v > 0.f ? v : 0.f;
    if (res) v += res[(size_t)r * N + n];
    out[(size_t)r * N + n] = v;
  }
}

// errc[0] = a-priori floor of the single-pass (fp16 operand) estimate's error in sqrt-cost space per unit of state
// RMS: ONE standard deviation of a first-order random-rounding model of the forward model:
//   one step adds, per state element, a residual error of RMS  u * (2/sqrt 3) * rho1 * rho2 * x_rms
//   (u = 2^-11 fp16 unit round-off; operands and weights of both layers rounded; rho1 = RMS row 2-norm of the STATE
//   columns of fwd.fc1, rho2 = max row 2-norm of fwd.fc2), T steps add in quadrature, sqrt(cost) = |s_T - goal| / sqrt(E)
//   sees it / sqrt(E).  It only guards against a degenerate sample (all 128 observed deviations ~0 by coincidence);
//   the generator's share, the action path and the gain of the dynamics are left to the data-derived term.
__global__ void __launch_bounds__(1024)
error_floor_kernel(const float* __restrict__ W1, const float* __restrict__ W2, int E, int H, int H_user, int T,
                   float* __restrict__ errc) {   // H: padded width of the copies; the RMS is over the caller's H_user rows
  __shared__ float s_sum[32], s_max[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float sum1 = 0.f, max2 = 0.f;
  for (size_t i = tid; i < (size_t)H * E; i += 1024) {   // state columns [0,E) of fwd.fc1 [H, E+4]
    const float v = W1[(i / E) * (size_t)(E + 4) + i % E];
    sum1 = fmaf(v, v, sum1);
  }
  for (int n = warp; n < E; n += 32) {                   // one warp per row of fwd.fc2 [E, H]
    float r = 0.f;
    for (int k = lane; k < H; k += 32) { const float v = W2[(size_t)n * H + k]; r = fmaf(v, v, r); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    max2 = fmaxf(max2, r);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum1 += __shfl_xor_sync(0xffffffffu, sum1, o);
  if (lane == 0) { s_sum[warp] = sum1; s_max[warp] = max2; }
  __syncthreads();
  if (tid == 0) {
    float a = 0.f, m = 0.f;
    for (int w = 0; w < 32; ++w) { a += s_sum[w]; m = fmaxf(m, s_max[w]); }
    const float rho1 = sqrtf(a / (float)H_user), rho2 = sqrtf(m);
    const float u = 1.f / 2048.f;
    errc[0] = u * 1.1547005f * rho1 * rho2 * sqrtf((float)T / (float)E);
  }
}

}  // namespace ndp

using namespace ndp;

extern "C" int ndp_abi_version(void) { return NDP_ABI_VERSION; }
extern "C" const char* ndp_last_error(void) { return g_err; }
extern "C" int ndp_device_check(void) { return device_check(); }
extern "C" uint64_t ndp_launch_count(void) { return g_launches.load(); }

extern "C" int ndp_planner_create(ndp_planner** out, const ndp_planner_desc* d) {
  NDP_CHECK_ARG(out && d, "planner_create: null argument");
  NDP_CHECK_ARG(d->E == 128 || d->E == 256, "planner_create: E must be 128 (the reference: the generator consumes [state 128 | goal 128 | noise], gan.py:67) or 256 (got %d)", d->E);
  NDP_CHECK_ARG(d->A == 4, "planner_create: A must be 4 (gan.py:71)");
  NDP_CHECK_ARG(d->nd >= 0 && d->nd <= 64, "planner_create: noise_dim must be in [0,64]");
  NDP_CHECK_ARG(d->H >= 1 && d->H <= 8192, "planner_create: H must be in [1,8192]");
  NDP_CHECK_ARG(d->T >= 1 && d->T <= 4096, "planner_create: T must be in [1,4096]");
  NDP_CHECK_ARG(d->precision >= NDP_PREC_FP32 && d->precision <= NDP_PREC_F16_REFINE, "planner_create: unknown precision mode %d", d->precision);
  if (d->precision != NDP_PREC_FP32)
    NDP_CHECK_ARG(d->E == 128 && d->nd <= 15 && d->H <= 2048, "planner_create: tensor-core modes need E=128, noise_dim<=15 and H<=2048 (got E=%d nd=%d H=%d); use NDP_PREC_FP32", d->E, d->nd, d->H);
  if (int rc = device_check()) return rc;
  ndp_planner* p = new ndp_planner();
  memset(p, 0, sizeof(*p));
  p->d = *d;
  // Any hidden width: the forward model is padded to the next multiple of 128 hidden units with zero fc1 rows, zero
  // fc1 bias and zero fc2 columns at load.  A padded unit is relu(0) = 0 and multiplies a zero weight: every sum gets
  // exact +0 terms, results are those of the caller's model (forward_encoder.py:107-114 has no width constraint).
  p->H_user = d->H;
  p->d.H = (d->H + 127) / 128 * 128;
  const int E = d->E, H = p->d.H, nd = d->nd;
  // 8 KB chunks of the fp32 weight stream per step: fc1 state columns E*128, fc2, fc3, fc4 (two halves); per hidden
  // chunk fwd.fc1 E*128 and fwd.fc2 128*E
  p->nq = (E / 16 + 4 + 4 + 16) + (E / 8) * (H / 128);
  const size_t sizes[] = {
      (size_t)p->nq * 2048,      // wstream
      (size_t)128 * E,           // w1g
      128, (size_t)(nd > 0 ? nd : 1) * 128, 64, 128, 256, 4 * 256, 4, (size_t)4 * H, (size_t)H, (size_t)E,
      64};                       // errc
  size_t total = 0;
  for (size_t s : sizes) total += (s + 63) / 64 * 64;
  // torch-layout copies for single-call APIs
  const size_t wt[7] = {(size_t)128 * (2 * E + nd), 64 * 128, 128 * 64, 256 * 128, 4 * 256, (size_t)H * (E + 4), (size_t)E * H};
  const size_t bt[7] = {128, 64, 128, 256, 4, (size_t)H, (size_t)E};
  for (int i = 0; i < 7; ++i) total += (wt[i] + 63) / 64 * 64 + (bt[i] + 63) / 64 * 64;
  cudaError_t e = cudaMalloc(&p->blob, total * sizeof(float));
  if (e != cudaSuccess) {
    set_error("planner_create: cudaMalloc(%zu) failed: %s", total * sizeof(float), cudaGetErrorString(e));
    delete p;
    return NDP_ERR_CUDA;
  }
  p->blob_floats = total;
  float* cur = p->blob;
  auto take = [&](size_t n) { float* r = cur; cur += (n + 63) / 64 * 64; return r; };
  p->wstream = take(sizes[0]); p->w1g = take(sizes[1]); p->b1 = take(sizes[2]); p->w1n = take(sizes[3]);
  p->b2 = take(sizes[4]); p->b3 = take(sizes[5]); p->b4 = take(sizes[6]); p->w5 = take(sizes[7]);
  p->b5 = take(sizes[8]); p->wf1a = take(sizes[9]); p->bf1 = take(sizes[10]); p->bf2 = take(sizes[11]);
  p->errc = take(sizes[12]);
  for (int i = 0; i < 7; ++i) { p->Wt[i] = take(wt[i]); p->bt[i] = take(bt[i]); }
  *out = p;
  return NDP_OK;
}

extern "C" void ndp_planner_destroy(ndp_planner* p) {
  if (!p) return;
  if (p->blob) cudaFree(p->blob);
  if (p->scratch) cudaFree(p->scratch);
  if (p->tc) tc_plan_destroy(p->tc);
  if (p->tc_refine) tc_plan_destroy(p->tc_refine);
  delete p;
}

static int pack(cudaStream_t s, float* dst, const float* src, int ld, int n0, int k0, int Kc, int N) {
  const int total = Kc * N;
  pack_T_kernel<<<(total + 255) / 256, 256, 0, s>>>(dst, src, ld, n0, k0, Kc, N);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_planner_load(ndp_planner* p, const float* const* W, const float* const* b,
                                int n_tensors, void* stream) {
  NDP_CHECK_ARG(p && W && b && n_tensors == 7, "planner_load: expects 7 weight and 7 bias tensors");
  for (int i = 0; i < 7; ++i) NDP_CHECK_ARG(W[i] && b[i], "planner_load: tensor %d is null", i);
  if (int rc = device_check()) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int E = p->d.E, H = p->d.H, Hu = p->H_user, nd = p->d.nd, C1 = 2 * E + nd;
  // caller's tensors have the caller's hidden width Hu; the internal copies are padded to H (zeros)
  const size_t wt[7] = {(size_t)128 * C1, 64 * 128, 128 * 64, 256 * 128, 4 * 256, (size_t)Hu * (E + 4), (size_t)E * Hu};
  const size_t bt[7] = {128, 64, 128, 256, 4, (size_t)Hu, (size_t)E};
  for (int i = 0; i < 7; ++i) {
    if (i == 6 && Hu != H) {                               // fc2 [E, Hu] -> [E, H]: pitched copy, zero tail columns
      NDP_CUDA(cudaMemsetAsync(p->Wt[6], 0, (size_t)E * H * sizeof(float), s));
      NDP_CUDA(cudaMemcpy2DAsync(p->Wt[6], (size_t)H * sizeof(float), W[6], (size_t)Hu * sizeof(float),
                                 (size_t)Hu * sizeof(float), E, cudaMemcpyDeviceToDevice, s));
    } else {
      if (i == 5 && Hu != H) NDP_CUDA(cudaMemsetAsync(p->Wt[5], 0, (size_t)H * (E + 4) * sizeof(float), s));   // zero tail rows
      NDP_CUDA(cudaMemcpyAsync(p->Wt[i], W[i], wt[i] * sizeof(float), cudaMemcpyDeviceToDevice, s));
    }
    if (i == 5 && Hu != H) NDP_CUDA(cudaMemsetAsync(p->bt[5], 0, (size_t)H * sizeof(float), s));
    NDP_CUDA(cudaMemcpyAsync(p->bt[i], b[i], bt[i] * sizeof(float), cudaMemcpyDeviceToDevice, s));
  }
  // ---- weight stream in consumption order (see k1_rollout_f32.cu), all blocks are W^T [K][N] row-major
  float* ws = p->wstream;
  int rc;
  if ((rc = pack(s, ws, p->Wt[0], C1, 0, 0, E, 128))) return rc;    ws += E * 128;     // gen.fc1 state cols
  if ((rc = pack(s, ws, p->Wt[1], 128, 0, 0, 128, 64))) return rc;  ws += 128 * 64;    // gen.fc2
  if ((rc = pack(s, ws, p->Wt[2], 64, 0, 0, 64, 128))) return rc;   ws += 64 * 128;    // gen.fc3
  for (int h = 0; h < 2; ++h) {                                                       // gen.fc4 halves
    if ((rc = pack(s, ws, p->Wt[3], 128, h * 128, 0, 128, 128))) return rc;
    ws += 128 * 128;
  }
  for (int c = 0; c < H / 128; ++c) {
    if ((rc = pack(s, ws, p->Wt[5], E + 4, c * 128, 0, E, 128))) return rc;    ws += E * 128;    // fwd.fc1 state cols, hidden chunk c
    for (int nh = 0; nh < E / 128; ++nh) {                                                       // fwd.fc2 rows of hidden chunk c, outputs in halves of 128
      if ((rc = pack(s, ws, p->Wt[6], H, nh * 128, c * 128, 128, 128))) return rc;
      ws += 128 * 128;
    }
  }
  // ---- side parameters
  NDP_CUDA(cudaMemcpy2DAsync(p->w1g, E * sizeof(float), p->Wt[0] + E, C1 * sizeof(float), E * sizeof(float), 128,
                             cudaMemcpyDeviceToDevice, s));                       // goal columns [128][E]
  if (nd > 0 && (rc = pack(s, p->w1n, p->Wt[0], C1, 0, 2 * E, nd, 128))) return rc;  // noise cols -> [nd][128]
  if ((rc = pack(s, p->wf1a, p->Wt[5], E + 4, 0, E, 4, H))) return rc;               // action cols -> [4][H]
  NDP_CUDA(cudaMemcpyAsync(p->b1, p->bt[0], 128 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->b2, p->bt[1], 64 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->b3, p->bt[2], 128 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->b4, p->bt[3], 256 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->w5, p->Wt[4], 4 * 256 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->b5, p->bt[4], 4 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->bf1, p->bt[5], (size_t)H * sizeof(float), cudaMemcpyDeviceToDevice, s));
  NDP_CUDA(cudaMemcpyAsync(p->bf2, p->bt[6], (size_t)E * sizeof(float), cudaMemcpyDeviceToDevice, s));
  error_floor_kernel<<<1, 1024, 0, s>>>(p->Wt[5], p->Wt[6], E, H, Hu, p->d.T, p->errc);
  NDP_LAUNCH_CHECK();
  if (p->d.precision != NDP_PREC_FP32) {
    if (p->tc) { tc_plan_destroy(p->tc); p->tc = nullptr; }
    if (p->tc_refine) { tc_plan_destroy(p->tc_refine); p->tc_refine = nullptr; }
    const int prec = p->d.precision;
    const int np = (prec == NDP_PREC_F16 || prec == NDP_PREC_F16_REFINE) ? 1 : (prec == NDP_PREC_F16X2 ? 2 : 3);
    if (int rc2 = tc_plan_create(&p->tc, np, E, nd, H, p->Wt, p->bt, s)) return rc2;
    if (prec == NDP_PREC_F16_REFINE)
      if (int rc2 = tc_plan_create(&p->tc_refine, 3, E, nd, H, p->Wt, p->bt, s)) return rc2;
  }
  p->loaded = true;
  return NDP_OK;
}

extern "C" int ndp_planner_workspace_bytes(const ndp_planner* p, int G, int K, size_t* bytes) {
  NDP_CHECK_ARG(p && bytes && G >= 0 && K >= 0, "planner_workspace_bytes: bad arguments");
  *bytes = (size_t)(G > 0 ? G : 1) * 128 * sizeof(float) + 256;
  if (p->tc) { const size_t t = tc_workspace_bytes(p->tc, G, K); if (t > *bytes) *bytes = t; }
  if (p->tc_refine) { const size_t t = tc_workspace_bytes(p->tc_refine, G, K); if (t > *bytes) *bytes = t; }
  return NDP_OK;
}

// balance: split the pairs of the two-tile kernel's last waves in time (k1_rollout_tc.cu: t2_split).  It shortens a
// rollout that has the GPU to itself (3 % at cfg2) and costs as much when the tail would have been filled by the next
// plan's CTAs anyway, so the pipelined entry (ndp_plan_select_bulk, PlanRing) leaves it off.
static int rollout_score_impl(ndp_planner* p, const float* e0, const float* goal, const float* noise,
                              int G, int K, float* cost, float* a0, float* state_T,
                              void* ws, size_t ws_bytes, void* stream, bool balance) {
  NDP_CHECK_ARG(p && p->loaded, "rollout_score: planner has no weights (call ndp_planner_load)");
  NDP_CHECK_ARG(G >= 0 && K >= 0, "rollout_score: bad G/K");
  if (int rc = device_check()) return rc;
  if (G == 0 || K == 0) return NDP_OK;
  NDP_CHECK_ARG(e0 && goal && cost && a0, "rollout_score: null tensor");
  NDP_CHECK_ARG(noise || p->d.nd == 0, "rollout_score: noise is required (noise_dim=%d)", p->d.nd);
  if (p->tc)
    return launch_rollout_tc(p->tc, e0, goal, noise, G, K, p->d.T, cost, a0, state_T, ws, ws_bytes, (cudaStream_t)stream,
                             nullptr, 0, nullptr, nullptr, balance);
  const size_t need = (size_t)G * 128 * sizeof(float);
  if (!ws || ws_bytes < need) { set_error("rollout_score: workspace %zu < %zu", ws_bytes, need); return NDP_ERR_WORKSPACE; }
  const int tile = rollout_f32_tile(p->d.E);
  const long long tiles = (long long)G * ((K + tile - 1) / tile);
  NDP_CHECK_ARG(tiles <= 0x7FFFFFFF, "rollout_score: G*K too large for one launch");
  cudaStream_t s = (cudaStream_t)stream;
  float* b1g = reinterpret_cast<float*>(ws);
  goal_bias_kernel<<<dim3(G, 16), 256, 0, s>>>(p->w1g, p->b1, goal, p->d.E, b1g);
  NDP_LAUNCH_CHECK();
  RolloutParams rp;
  rp.wstream = p->wstream; rp.b1g = b1g; rp.w1n = p->w1n; rp.b2 = p->b2; rp.b3 = p->b3; rp.b4 = p->b4;
  rp.w5 = p->w5; rp.b5 = p->b5; rp.wf1a = p->wf1a; rp.bf1 = p->bf1; rp.bf2 = p->bf2;
  rp.e0 = e0; rp.goal = goal; rp.noise = noise; rp.cost = cost; rp.a0 = a0; rp.state_T = state_T;
  rp.G = G; rp.K = K; rp.T = p->d.T; rp.nd = p->d.nd; rp.H = p->d.H; rp.E = p->d.E;
  rp.tiles_per_goal = (K + tile - 1) / tile; rp.nq = p->nq;
  return launch_rollout_f32(rp, s);
}

extern "C" int ndp_rollout_score(ndp_planner* p, const float* e0, const float* goal, const float* noise,
                                 int G, int K, float* cost, float* a0, float* state_T,
                                 void* ws, size_t ws_bytes, void* stream) {
  return rollout_score_impl(p, e0, goal, noise, G, K, cost, a0, state_T, ws, ws_bytes, stream, true);
}

// handle-owned activation scratch of the single-call entry points (grown on demand).  One planner handle must not be
// used by two streams at once through these calls; growing is refused while the stream is being captured (warm the
// call up eagerly with the largest batch first, as ClosedLoopSweep.capture does).
static int ensure_scratch(ndp_planner* p, size_t floats, cudaStream_t s) {
  if (p->scratch_floats >= floats) return NDP_OK;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone) {
    set_error("generator_forward / forward_step: scratch must grow to %zu floats while the stream is capturing; run the call eagerly once first", floats);
    return NDP_ERR_WORKSPACE;
  }
  if (p->scratch) cudaFree(p->scratch);
  p->scratch = nullptr; p->scratch_floats = 0;
  NDP_CUDA(cudaMalloc(&p->scratch, floats * sizeof(float)));
  p->scratch_floats = floats;
  return NDP_OK;
}

static int linear(cudaStream_t s, const float* in0, int K0, const float* in1, int K1, const float* W,
                  const float* b, int N, int B, int relu, const float* res, float* out) {
  const long long warps = (long long)B * N;
  const long long blocks = (warps + 7) / 8;
  NDP_CHECK_ARG(blocks <= 0x7FFFFFFF, "batch too large");
  linear_kernel<<<(unsigned)blocks, 256, 0, s>>>(in0, K0, in1, K1, W, b, N, B, relu, res, out);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_generator_forward(ndp_planner* p, const float* z, int B, float* a, void* stream) {
  NDP_CHECK_ARG(p && p->loaded, "generator_forward: planner has no weights");
  NDP_CHECK_ARG(B >= 0, "generator_forward: bad B");
  if (int rc = device_check()) return rc;
  if (B == 0) return NDP_OK;
  NDP_CHECK_ARG(z && a, "generator_forward: null tensor");
  cudaStream_t s = (cudaStream_t)stream;
  if (int rc = ensure_scratch(p, (size_t)B * 512, s)) return rc;
  float* h1 = p->scratch; float* h2 = h1 + (size_t)B * 256;
  const int C1 = 2 * p->d.E + p->d.nd;
  int rc;
  if ((rc = linear(s, z, C1, nullptr, 0, p->Wt[0], p->bt[0], 128, B, 1, nullptr, h1))) return rc;
  if ((rc = linear(s, h1, 128, nullptr, 0, p->Wt[1], p->bt[1], 64, B, 1, nullptr, h2))) return rc;
  if ((rc = linear(s, h2, 64, nullptr, 0, p->Wt[2], p->bt[2], 128, B, 1, nullptr, h1))) return rc;
  if ((rc = linear(s, h1, 128, nullptr, 0, p->Wt[3], p->bt[3], 256, B, 1, nullptr, h2))) return rc;
  return linear(s, h2, 256, nullptr, 0, p->Wt[4], p->bt[4], 4, B, 0, nullptr, a);
}

extern "C" int ndp_forward_step(ndp_planner* p, const float* state, const float* action, int B,
                                float* next_state, void* stream) {
  NDP_CHECK_ARG(p && p->loaded, "forward_step: planner has no weights");
  NDP_CHECK_ARG(B >= 0, "forward_step: bad B");
  if (int rc = device_check()) return rc;
  if (B == 0) return NDP_OK;
  NDP_CHECK_ARG(state && action && next_state, "forward_step: null tensor");
  cudaStream_t s = (cudaStream_t)stream;
  if (int rc = ensure_scratch(p, (size_t)B * (size_t)(p->d.H > 512 ? p->d.H : 512), s)) return rc;
  int rc;
  if ((rc = linear(s, state, p->d.E, action, 4, p->Wt[5], p->bt[5], p->d.H, B, 1, nullptr, p->scratch))) return rc;
  return linear(s, p->scratch, p->d.H, nullptr, 0, p->Wt[6], p->bt[6], p->d.E, B, 0, state, next_state);
}

// ---- one-call plan ---------------------------------------------------------------------------
static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

struct PlanWs {
  size_t rollout, select, list, est, packed, margin, total;
};
static int plan_ws_layout(const ndp_planner* p, int G, int K, int topk, PlanWs* w) {
  size_t r = 0, sel = 0;
  if (int rc = ndp_planner_workspace_bytes(p, G, K, &r)) return rc;
  const int tk = p->tc_refine ? NDP_REFINE_R : topk;
  if (int rc = ndp_select_workspace_bytes(G, K, tk, &sel)) return rc;
  w->rollout = align256(r);
  w->select = align256(sel);
  const size_t GR = (size_t)(G > 0 ? G : 1) * NDP_REFINE_R;
  w->list = p->tc_refine ? align256(GR * sizeof(int)) : 0;
  w->est = p->tc_refine ? align256(GR * sizeof(float)) : 0;
  w->packed = p->tc_refine ? align256((GR + (size_t)(G > 0 ? G : 1)) * sizeof(unsigned long long)) : 0;
  w->margin = align256((size_t)(G > 0 ? G : 1) * 2 * sizeof(float));
  w->total = w->rollout + w->select + w->list + w->est + w->packed + w->margin;
  return NDP_OK;
}

extern "C" int ndp_plan_workspace_bytes(const ndp_planner* p, int G, int K, int topk, size_t* bytes) {
  NDP_CHECK_ARG(p && bytes && G >= 0 && K >= 0 && topk >= 0 && topk <= 128, "plan_workspace_bytes: bad arguments");
  PlanWs w;
  if (int rc = plan_ws_layout(p, G, K, topk, &w)) return rc;
  *bytes = w.total;
  return NDP_OK;
}

// phase bit 0: bulk rollout (+ the whole selection when not refining)
// phase bit 1: selection of the refine list + refine pass + selection from refined costs + verification
static int plan_select_impl(int phase, ndp_planner* p, const float* e0, const float* goal, const float* noise, int G,
                            int K, int topk, int64_t idx_base, float eps, float* cost, float* a0,
                            uint64_t* packed_min, int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws,
                            size_t ws_bytes, void* stream) {
  NDP_CHECK_ARG(p && p->loaded, "plan_select: planner has no weights (call ndp_planner_load)");
  NDP_CHECK_ARG(G >= 0 && K >= 0 && topk >= 0 && topk <= 128, "plan_select: bad G/K/topk");
  NDP_CHECK_ARG(!p->tc_refine || topk <= 64, "plan_select: topk must be <= 64 in refine mode (got %d)", topk);
  NDP_CHECK_ARG(topk == 0 || (topk_idx && topk_cost), "plan_select: topk_idx/topk_cost required when topk > 0");
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  NDP_CHECK_ARG(packed_min, "plan_select: null packed_min");
  PlanWs w;
  if (int rc = plan_ws_layout(p, G, K, topk, &w)) return rc;
  if (!ws || ws_bytes < w.total) { set_error("plan_select: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  NDP_CHECK_ARG(((uintptr_t)ws & 255) == 0, "plan_select: workspace must be 256-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  char* base = reinterpret_cast<char*>(ws);
  void* ws_roll = base;
  void* ws_sel = base + w.rollout;
  const int R = NDP_REFINE_R;
  int* list = reinterpret_cast<int*>(base + w.rollout + w.select);
  float* est = reinterpret_cast<float*>(base + w.rollout + w.select + w.list);
  unsigned long long* packed_c = reinterpret_cast<unsigned long long*>(base + w.rollout + w.select + w.list + w.est);
  unsigned long long* pm_est = packed_c + (size_t)G * R;
  float* margin = reinterpret_cast<float*>(base + w.rollout + w.select + w.list + w.est + w.packed);
  if (phase & 1) {
    if (int rc = rollout_score_impl(p, e0, goal, noise, G, K, cost, a0, nullptr, ws_roll, w.rollout, stream,
                                    /*balance=*/phase != 1)) return rc;
    if (!p->tc_refine) {
      if (int rc = ndp_select(cost, G, K, topk, idx_base, packed_min, topk_idx, topk_cost, ws_sel, w.select, stream)) return rc;
      if (verified)   // K=0 <= R marks every goal verified: nothing was estimated
        if (int rc = verify_refine(nullptr, nullptr, topk_cost, reinterpret_cast<const unsigned long long*>(packed_min),
                                   nullptr, p->d.E, p->errc, 0.f, G, 0, R, topk, verified, margin, s)) return rc;
      return NDP_OK;
    }
  }
  if ((phase & 4) && p->tc_refine && verified && K > R) {
    // Exact fallback: goals whose refined selection could NOT be proven (verified[g] == 0) are re-rolled over ALL K
    // candidates with split operands and selected from those costs; decided on the device (skip mask = verified),
    // no host round trip.  CTAs / blocks of verified goals exit immediately.  1 = proven by the bound, 2 = re-rolled.
    if (int rc = launch_rollout_tc(p->tc_refine, e0, goal, noise, G, K, p->d.T, cost, a0, nullptr, ws_roll, w.rollout,
                                   s, nullptr, 0, nullptr, verified)) return rc;
    if (int rc = select_masked(cost, G, K, topk, (long long)idx_base, reinterpret_cast<unsigned long long*>(packed_min),
                               topk_idx, topk_cost, ws_sel, w.select, verified, s)) return rc;
    if (int rc = mark_fallback(verified, G, s)) return rc;
    return NDP_OK;
  }
  if ((phase & 2) && p->tc_refine) {
    // the R best ESTIMATES per goal (local indices, ascending; -1 padded when K < R)
    if (int rc = ndp_select(cost, G, K, R, 0, reinterpret_cast<uint64_t*>(pm_est), list, est, ws_sel, w.select, stream)) return rc;
    // re-roll them with split operands; cost / a0 entries are overwritten, packed keys go to packed_c.
    // (the per-goal bias block is rebuilt in ws_roll: the bulk pass of THIS plan has finished with it)
    if (int rc = launch_rollout_tc(p->tc_refine, e0, goal, noise, G, K, p->d.T, cost, a0, nullptr, ws_roll, w.rollout, s,
                                   list, R, packed_c)) return rc;
    // the plan from the refined costs (ties -> lower candidate index, like the reference scan)
    if (int rc = select_packed(packed_c, G, R, topk, (long long)idx_base,
                               reinterpret_cast<unsigned long long*>(packed_min), topk_idx, topk_cost, s)) return rc;
    if (verified) {
      if (int rc = verify_refine(est, packed_c, topk_cost, reinterpret_cast<const unsigned long long*>(packed_min), goal,
                                 p->d.E, p->errc, eps, G, K, R, topk, verified, margin, s)) return rc;
    }
  }
  return NDP_OK;
}

extern "C" int ndp_plan_select(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                               int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                               int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                               void* stream) {
  return plan_select_impl(3, p, e0, goal, noise, G, K, topk, idx_base, eps, cost, a0, packed_min, topk_idx, topk_cost,
                          verified, ws, ws_bytes, stream);
}
extern "C" int ndp_plan_fallback(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                                 int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                                 int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                                 void* stream) {
  return plan_select_impl(4, p, e0, goal, noise, G, K, topk, idx_base, eps, cost, a0, packed_min, topk_idx, topk_cost,
                          verified, ws, ws_bytes, stream);
}
extern "C" int ndp_plan_select_bulk(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G,
                                    int K, int topk, int64_t idx_base, float eps, float* cost, float* a0,
                                    uint64_t* packed_min, int32_t* topk_idx, float* topk_cost, int32_t* verified,
                                    void* ws, size_t ws_bytes, void* stream) {
  return plan_select_impl(1, p, e0, goal, noise, G, K, topk, idx_base, eps, cost, a0, packed_min, topk_idx, topk_cost,
                          verified, ws, ws_bytes, stream);
}
extern "C" int ndp_plan_select_refine(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G,
                                      int K, int topk, int64_t idx_base, float eps, float* cost, float* a0,
                                      uint64_t* packed_min, int32_t* topk_idx, float* topk_cost, int32_t* verified,
                                      void* ws, size_t ws_bytes, void* stream) {
  return plan_select_impl(2, p, e0, goal, noise, G, K, topk, idx_base, eps, cost, a0, packed_min, topk_idx, topk_cost,
                          verified, ws, ws_bytes, stream);
}

extern "C" int ndp_plan_verify_info(const ndp_planner* p, int G, int K, int topk, const void* ws, size_t ws_bytes,
                                    float* info, void* stream) {
  NDP_CHECK_ARG(p && info && G >= 0 && K >= 0 && topk >= 0 && topk <= 128, "plan_verify_info: bad arguments");
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  PlanWs w;
  if (int rc = plan_ws_layout(p, G, K, topk, &w)) return rc;
  if (!ws || ws_bytes < w.total) { set_error("plan_verify_info: workspace %zu < %zu", ws_bytes, w.total); return NDP_ERR_WORKSPACE; }
  const char* src = reinterpret_cast<const char*>(ws) + w.rollout + w.select + w.list + w.est + w.packed;
  NDP_CUDA(cudaMemcpyAsync(info, src, (size_t)G * 2 * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return NDP_OK;
}

// ---- device sampler: U[0,1) noise of diverse_sampling (mpc_eval.py:78-83) drawn next to its
// consumer.  Counter-based Philox4x32-10 (Salmon et al., SC'11): value i is word (i & 3) of
// philox(counter = {i/4 lo, i/4 hi, offset lo, offset hi}, key = {seed lo, seed hi}), mapped to [0,1) by its top
// 24 bits -- reproducible from (seed, offset) alone, independent of launch geometry.
namespace ndp {
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
__global__ void sample_uniform_kernel(unsigned long long seed, unsigned long long offset, float* __restrict__ out,
                                      size_t n) {
  const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // one Philox block = 4 values
  if (q * 4 >= n) return;
  uint32_t c[4] = {(uint32_t)q, (uint32_t)(q >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; r++) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
#pragma unroll
  for (int j = 0; j < 4; j++)
    if (q * 4 + j < n) out[q * 4 + j] = (float)(c[j] >> 8) * (1.0f / 16777216.0f);
}
}  // namespace ndp

extern "C" int ndp_sample_uniform(uint64_t seed, uint64_t offset, float* out, size_t n, void* stream) {
  NDP_CHECK_ARG(out || n == 0, "ndp_sample_uniform: null output");
  if (int rc = ndp::device_check()) return rc;
  if (n == 0) return NDP_OK;
  const size_t q = (n + 3) / 4;
  ndp::sample_uniform_kernel<<<(unsigned)((q + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, offset, out, n);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
