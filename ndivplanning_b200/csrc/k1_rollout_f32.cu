// K1 (fp32 parity mode) — fused generator MLP + T-step latent rollout + terminal cost.
//
// One CTA owns a tile of M candidates of one goal for the WHOLE horizon: the state
// [M x E] never leaves shared memory between steps, hidden activations never touch HBM.
// E = 128 (the reference's latent width, M = 128, 256 threads) or E = 256 (the alternative reading of BASELINE cfg2's
// "embedding dim 256"; the state image doubles, so M = 64 with 128 threads keeps the CTA inside 227 KB).
// Replaces the reference loop mpc_eval.py:124-157 (+ :161 per-candidate MSE):
//   per step  z = [state | goal | noise]  -> gan.Decoder.forward (gan.py:79-86) -> action
//             state <- state + fc2(relu(fc1([state | action])))   (forward_encoder.py:105-114 contract)
//   after T   cost = mean_E (state - goal)^2                        (mpc_eval.py:161)
// The goal half of generator fc1 is constant per plan and folded into a per-goal bias;
// noise (nd) and action (4) columns are rank-nd / rank-4 updates in the epilogues.
//
// All contractions are fp32 FFMA with 8x8 register tiles; weights stream from L2 through a
// 2-stage cp.async ring as a flat sequence of 8 KB chunks prepacked in consumption order.
// Activations live k-major ([feature][candidate], stride 132) so both fragments are LDS.128.
#include "planner.cuh"

namespace ndp {

static constexpr int kChunk = 2048;     // floats per weight chunk (8 KB)
// M candidates per CTA, 2 M threads (16 column groups x M/8 row groups of 8x8 register tiles), activation stride M + 4

__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

template <int M>
struct Stream {
  const float* base;   // packed weights of one step
  float* ring;         // 2 x kChunk floats in smem
  int q;               // next chunk to consume (within the step)
  int nq;              // chunks per step
  __device__ __forceinline__ void prefetch(int chunk) {
    const float* src = base + (size_t)chunk * kChunk;
    float* dst = ring + (chunk & 1) * kChunk;
#pragma unroll
    for (int i = 0; i < kChunk / 4 / (2 * M); ++i)
      cp_async16(dst + (threadIdx.x + i * 2 * M) * 4, src + (threadIdx.x + i * 2 * M) * 4);
    cp_async_commit();
  }
};

// acc[8][NT] += X[M x K] * Wt[K x N]; X k-major in smem; N = 16*NT; consumes K*N/kChunk chunks.
// NOTE: chunk parity (q & 1) must stay consistent with prefetch(): nq is even.
template <int NT, int M>
__device__ __forceinline__ void gemm_stream(float (&acc)[8][NT], const float* __restrict__ Xs, int K, Stream<M>& st) {
  constexpr int N = 16 * NT;
  constexpr int kLd = M + 4;
  constexpr int KPC = kChunk / N;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int k0 = 0; k0 < K; k0 += KPC) {
    cp_async_wait_all();
    __syncthreads();
    int nxt = st.q + 1;
    if (nxt == st.nq) nxt = 0;
    st.prefetch(nxt);
    const float* Wc = st.ring + (st.q & 1) * kChunk;
    const float* Xk = Xs + (size_t)k0 * kLd + ty * 8;
#pragma unroll
    for (int kk = 0; kk < KPC; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(Xk + kk * kLd);
      const float4 a1 = *reinterpret_cast<const float4*>(Xk + kk * kLd + 4);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[NT];
      const float4 b0 = *reinterpret_cast<const float4*>(Wc + kk * N + tx * 4);
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
      if (NT == 8) {
        const float4 b1 = *reinterpret_cast<const float4*>(Wc + kk * N + 64 + tx * 4);
        b[NT - 4] = b1.x; b[NT - 3] = b1.y; b[NT - 2] = b1.z; b[NT - 1] = b1.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    st.q = nxt;
  }
}

template <int NT>
__device__ __forceinline__ int col_of(int j) {
  const int tx = threadIdx.x & 15;
  return (NT == 8 && j >= 4) ? 64 + tx * 4 + (j - 4) : tx * 4 + j;
}

template <int NT>
__device__ __forceinline__ void zero_acc(float (&acc)[8][NT]) {
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) acc[i][j] = 0.f;
}

// Out[n][m] = relu(acc[m][n])  (k-major for the next layer)
template <int NT, int M>
__device__ __forceinline__ void store_relu(const float (&acc)[8][NT], float* __restrict__ Out) {
  constexpr int kLd = M + 4;
  const int ty = threadIdx.x >> 4;
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    float* o = Out + (size_t)col_of<NT>(j) * kLd + ty * 8;
    *reinterpret_cast<float4*>(o) = make_float4(fmaxf(acc[0][j], 0.f), fmaxf(acc[1][j], 0.f),
                                                fmaxf(acc[2][j], 0.f), fmaxf(acc[3][j], 0.f));
    *reinterpret_cast<float4*>(o + 4) = make_float4(fmaxf(acc[4][j], 0.f), fmaxf(acc[5][j], 0.f),
                                                    fmaxf(acc[6][j], 0.f), fmaxf(acc[7][j], 0.f));
  }
}

template <int EC>
__global__ void __launch_bounds__(EC == 1 ? 256 : 128, 1) rollout_f32_kernel(const RolloutParams p) {
  constexpr int kM = EC == 1 ? 128 : 64, kLd = kM + 4, kRThreads = 2 * kM, E = 128 * EC;
  extern __shared__ __align__(16) float smem[];
  float* S = smem;                       // state        [E][kLd]
  float* bufA = S + E * kLd;             // activations  [128][kLd]
  float* bufB = bufA + 128 * kLd;        // activations  [128][kLd]
  float* ring = bufB + 128 * kLd;        // weight ring  2 x kChunk
  float* s_act = ring + 2 * kChunk;      // action       [4][kLd]

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int g = blockIdx.x / p.tiles_per_goal;
  const int m0 = (blockIdx.x % p.tiles_per_goal) * kM;

  Stream<kM> st{p.wstream, ring, 0, p.nq};
  st.prefetch(0);
  for (int idx = tid; idx < E * kM; idx += kRThreads) {
    const int e = idx / kM, m = idx % kM;
    S[e * kLd + m] = __ldg(p.e0 + (size_t)g * E + e);
  }

  float acc[8][8];
  for (int t = 0; t < p.T; ++t) {
    // ---------------- generator fc1: [state | goal | noise] -> 128   (gan.py:81)
    zero_acc<8>(acc);
    gemm_stream<8, kM>(acc, S, E, st);
    {
      float nz[8][4];
      const float* nbase = p.noise + (((size_t)t * p.G + g) * p.K) * p.nd;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = col_of<8>(j);
        const float bias = __ldg(p.b1g + (size_t)g * 128 + n);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i][j] += bias;
      }
      for (int d0 = 0; d0 < p.nd; d0 += 4) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int cand = m0 + ty * 8 + i;
#pragma unroll
          for (int dd = 0; dd < 4; ++dd)
            nz[i][dd] = (cand < p.K && d0 + dd < p.nd) ? __ldg(nbase + (size_t)cand * p.nd + d0 + dd) : 0.f;
        }
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) {
          if (d0 + dd < p.nd) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float w = __ldg(p.w1n + (size_t)(d0 + dd) * 128 + col_of<8>(j));
#pragma unroll
              for (int i = 0; i < 8; ++i) acc[i][j] = fmaf(nz[i][dd], w, acc[i][j]);
            }
          }
        }
      }
      store_relu<8, kM>(acc, bufA);
    }
    // ---------------- fc2: 128 -> 64   (gan.py:82)
    {
      float acc4[8][4];
      zero_acc<4>(acc4);
      gemm_stream<4, kM>(acc4, bufA, 128, st);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float bias = __ldg(p.b2 + col_of<4>(j));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc4[i][j] += bias;
      }
      store_relu<4, kM>(acc4, bufB);
    }
    // ---------------- fc3: 64 -> 128   (gan.py:83)
    zero_acc<8>(acc);
    gemm_stream<8, kM>(acc, bufB, 64, st);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float bias = __ldg(p.b3 + col_of<8>(j));
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i][j] += bias;
    }
    store_relu<8, kM>(acc, bufA);
    // ---------------- fc4: 128 -> 256 in two halves, fc5 (256 -> 4) folded in   (gan.py:84-85)
    const int am = tid % kM, ap = (tid / kM) * 2;   // this thread's (candidate, action pair)
    float av0 = 0.f, av1 = 0.f;
    for (int h = 0; h < 2; ++h) {
      zero_acc<8>(acc);
      gemm_stream<8, kM>(acc, bufA, 128, st);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float bias = __ldg(p.b4 + h * 128 + col_of<8>(j));
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i][j] += bias;
      }
      store_relu<8, kM>(acc, bufB);
      __syncthreads();
      const float* w5a = p.w5 + (size_t)ap * 256 + h * 128;
      const float* w5b = w5a + 256;
#pragma unroll 8
      for (int n = 0; n < 128; ++n) {
        const float hv = bufB[n * kLd + am];
        av0 = fmaf(hv, __ldg(w5a + n), av0);
        av1 = fmaf(hv, __ldg(w5b + n), av1);
      }
    }
    av0 += __ldg(p.b5 + ap);
    av1 += __ldg(p.b5 + ap + 1);
    s_act[ap * kLd + am] = av0;
    s_act[(ap + 1) * kLd + am] = av1;
    if (t == 0 && m0 + am < p.K) {   // step-0 actions are the ones executed (mpc_eval.py:150-151)
      float* a0p = p.a0 + ((size_t)g * p.K + m0 + am) * 4 + ap;
      a0p[0] = av0;
      a0p[1] = av1;
    }
    // ---------------- forward model: state += fc2(relu(fc1([state | action])))
    float acc2[EC][8][8];                    // residual accumulators: E outputs in halves of 128
#pragma unroll
    for (int nh = 0; nh < EC; ++nh) zero_acc<8>(acc2[nh]);
    for (int c = 0; c < p.H / 128; ++c) {
      zero_acc<8>(acc);
      gemm_stream<8, kM>(acc, S, E, st);   // first chunk barrier also publishes s_act
      {
        float a4[8][4];
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          const float4 v0 = *reinterpret_cast<const float4*>(s_act + d * kLd + ty * 8);
          const float4 v1 = *reinterpret_cast<const float4*>(s_act + d * kLd + ty * 8 + 4);
          a4[0][d] = v0.x; a4[1][d] = v0.y; a4[2][d] = v0.z; a4[3][d] = v0.w;
          a4[4][d] = v1.x; a4[5][d] = v1.y; a4[6][d] = v1.z; a4[7][d] = v1.w;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int n = c * 128 + col_of<8>(j);
          const float bias = __ldg(p.bf1 + n);
          const float w0 = __ldg(p.wf1a + n), w1 = __ldg(p.wf1a + p.H + n);
          const float w2 = __ldg(p.wf1a + 2 * p.H + n), w3 = __ldg(p.wf1a + 3 * p.H + n);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float v = acc[i][j] + bias;
            v = fmaf(a4[i][0], w0, v); v = fmaf(a4[i][1], w1, v);
            v = fmaf(a4[i][2], w2, v); v = fmaf(a4[i][3], w3, v);
            acc[i][j] = v;
          }
        }
      }
      store_relu<8, kM>(acc, bufA);
#pragma unroll
      for (int nh = 0; nh < EC; ++nh) gemm_stream<8, kM>(acc2[nh], bufA, 128, st);
    }
    // ---------------- residual update (eval mode: state + residual, forward_encoder.py:113-114)
    const bool last = (t == p.T - 1);
    float csum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int nh = 0; nh < EC; ++nh) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = nh * 128 + col_of<8>(j);
        const float bias = __ldg(p.bf2 + n);
        float* sp = S + (size_t)n * kLd + ty * 8;
        const float gl = last ? __ldg(p.goal + (size_t)g * E + n) : 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float v = sp[i] + (acc2[nh][i][j] + bias);
          sp[i] = v;
          acc2[nh][i][j] = v;
          const float df = v - gl;
          csum[i] = fmaf(df, df, csum[i]);
        }
      }
    }
    if (last) {
      // terminal cost: mean over E of squared error (nn.MSELoss, mpc_eval.py:161)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float v = csum[i];
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        const int cand = m0 + ty * 8 + i;
        if (tx == 0 && cand < p.K) p.cost[(size_t)g * p.K + cand] = v * (1.0f / (float)E);
      }
      if (p.state_T) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int cand = m0 + ty * 8 + i;
          if (cand < p.K) {
            float* o = p.state_T + ((size_t)g * p.K + cand) * E;
#pragma unroll
            for (int nh = 0; nh < EC; ++nh)
#pragma unroll
              for (int j = 0; j < 8; ++j) o[nh * 128 + col_of<8>(j)] = acc2[nh][i][j];
          }
        }
      }
    }
  }
  cp_async_wait_all();
}

int rollout_f32_tile(int E) { return E == 128 ? 128 : 64; }
int rollout_f32_smem_bytes(int E) {
  const int ld = rollout_f32_tile(E) + 4;
  return ((E + 2 * 128) * ld + 2 * kChunk + 4 * ld) * (int)sizeof(float);
}

int launch_rollout_f32(const RolloutParams& p, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_devs{0};
  if (first_use_on_device(attr_devs)) {
    NDP_CUDA(cudaFuncSetAttribute(rollout_f32_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, rollout_f32_smem_bytes(128)));
    NDP_CUDA(cudaFuncSetAttribute(rollout_f32_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, rollout_f32_smem_bytes(256)));
  }
  const int grid = p.G * p.tiles_per_goal, smem = rollout_f32_smem_bytes(p.E);
  if (p.E == 128) rollout_f32_kernel<1><<<grid, 256, smem, s>>>(p);
  else if (p.E == 256) rollout_f32_kernel<2><<<grid, 128, smem, s>>>(p);
  else { set_error("rollout_f32: E must be 128 or 256 (got %d)", p.E); return NDP_ERR_ARG; }
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

}  // namespace ndp
