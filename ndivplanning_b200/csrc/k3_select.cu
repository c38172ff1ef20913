// K3 — segmented argmin / top-k over candidate costs.
// Replaces the per-candidate python scan of the reference (mpc_eval.py:122,159-165):
//   strict '<' from min_error = 1e10, first minimal index wins, NaN never wins.
// Every cost becomes a u64  (ordered_key(cost) << 32) | index  so that one unsigned min
// implements (cost, index) lexicographic order; the same word is what the multi-GPU
// min-allreduce exchanges.  HBM-bound: G*K*4 bytes read once, float4-vectorised.
#include <stdlib.h>
#include <mutex>
#include "common.cuh"
#include "peer.cuh"

namespace ndp {

static constexpr int kSelThreads = 256;
static constexpr int kSelTile = 1024;     // elements per block iteration (4 per thread)
static constexpr int kSelBuf = 2048;      // candidate buffer entries (u64) in smem
static constexpr unsigned long long kPadKey = ~0ull;

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t < v ? t : v;
  }
  return v;
}

// in-place ascending bitonic sort of buf[0..len) by the whole block (len = power of two <= kSelBuf)
__device__ void block_bitonic_sort(unsigned long long* buf, int len) {
  for (int k = 2; k <= len; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < len; i += kSelThreads) {
        int ixj = i ^ j;
        if (ixj > i) {
          unsigned long long a = buf[i], b = buf[ixj];
          bool up = ((i & k) == 0);
          if ((a > b) == up) { buf[i] = b; buf[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
}

struct TopkState {
  unsigned long long* buf;   // [kSelBuf]
  int* cnt;                  // pending count
  unsigned long long* thr;   // current admission threshold (exclusive)
};

// keep the `topk` smallest of buf[0..*cnt), set threshold; block-wide
__device__ void topk_compact(TopkState s, int topk) {
  __syncthreads();
  int n = *s.cnt;
  int len = 32;
  while (len < n) len <<= 1;
  for (int i = n + threadIdx.x; i < len; i += kSelThreads) s.buf[i] = kPadKey;
  __syncthreads();
  block_bitonic_sort(s.buf, len);
  if (threadIdx.x == 0) {
    int m = n < topk ? n : topk;
    *s.cnt = m;
    *s.thr = (m == topk) ? s.buf[topk - 1] : kPadKey;
  }
  __syncthreads();
}

__device__ __forceinline__ void topk_offer(TopkState s, unsigned long long p) {
  if (p < *s.thr) {
    int pos = atomicAdd(s.cnt, 1);
    s.buf[pos] = p;   // capacity guaranteed by the compaction schedule
  }
}

// MODE 0: input fp32 costs;  MODE 1: input packed u64 (merge of partials)
template <int MODE>
__global__ void __launch_bounds__(kSelThreads)
select_kernel(const float* __restrict__ cost, const unsigned long long* __restrict__ packed_in,
              int K, int chunk_len, int topk, int final_stage, long long idx_base,
              unsigned long long* __restrict__ out_min,     // [G, chunks] or [G]
              unsigned long long* __restrict__ out_topk,    // [G, chunks, topk] partial
              int* __restrict__ topk_idx, float* __restrict__ topk_cost,
              const int* __restrict__ skip) {          // skip[g] != 0: goal g keeps its previous results
  if (skip && skip[blockIdx.y]) return;
  __shared__ unsigned long long s_buf[kSelBuf];
  __shared__ unsigned long long s_warpmin[kSelThreads / 32];
  __shared__ unsigned long long s_thr;
  __shared__ int s_cnt;
  const int g = blockIdx.y, chunk = blockIdx.x, chunks = gridDim.x;
  const int begin = chunk * chunk_len;
  int end = begin + chunk_len; if (end > K) end = K;
  TopkState st{s_buf, &s_cnt, &s_thr};
  if (threadIdx.x == 0) { s_cnt = 0; s_thr = kPadKey; }
  __syncthreads();

  unsigned long long best = kPadKey;
  const size_t seg = (size_t)g * (size_t)K;
  for (int base = begin; base < end; base += kSelTile) {
    unsigned long long p[4];
    const int i0 = base + threadIdx.x * 4;
    if (MODE == 0) {
      const float* src = cost + seg + i0;
      float c[4];
      if (i0 + 3 < end && ((((size_t)(seg + i0)) & 3) == 0)) {
        float4 v = __ldg(reinterpret_cast<const float4*>(src));
        c[0] = v.x; c[1] = v.y; c[2] = v.z; c[3] = v.w;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          p[q] = ((unsigned long long)ordered_key(c[q] + 0.0f) << 32) | (unsigned)(i0 + q);
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          p[q] = (i0 + q < end)
                     ? (((unsigned long long)ordered_key(__ldg(src + q) + 0.0f) << 32) | (unsigned)(i0 + q))
                     : kPadKey;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) p[q] = (i0 + q < end) ? packed_in[seg + i0 + q] : kPadKey;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) best = p[q] < best ? p[q] : best;
    if (topk > 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (p[q] != kPadKey || MODE == 0) { if (i0 + q < end) topk_offer(st, p[q]); }
      __syncthreads();
      const int pending = s_cnt;   // uniform: read between two barriers
      __syncthreads();
      if (pending + kSelTile > kSelBuf) topk_compact(st, topk);
    }
  }
  // ---- argmin: warp shuffle -> smem -> thread 0
  best = warp_min_u64(best);
  if ((threadIdx.x & 31) == 0) s_warpmin[threadIdx.x >> 5] = best;
  if (topk > 0) topk_compact(st, topk); else __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long m = s_warpmin[0];
#pragma unroll
    for (int w = 1; w < kSelThreads / 32; ++w) m = s_warpmin[w] < m ? s_warpmin[w] : m;
    if (final_stage) {
      // reference rule: only costs strictly below 1e10 may win (mpc_eval.py:122,161-164)
      float c = key_to_float((uint32_t)(m >> 32));
      if (m == kPadKey || !(c < 1e10f)) m = kInvalidPacked;
      else m = (m & 0xFFFFFFFF00000000ull) | (unsigned long long)((long long)(m & 0xFFFFFFFFull) + idx_base);
      out_min[g] = m;
    } else {
      out_min[(size_t)g * chunks + chunk] = m;
    }
  }
  if (topk > 0) {
    const int n = s_cnt;
    for (int i = threadIdx.x; i < topk; i += kSelThreads) {
      unsigned long long v = i < n ? s_buf[i] : kPadKey;
      if (final_stage) {
        if (topk_idx) topk_idx[(size_t)g * topk + i] = (i >= n) ? -1 : (int)((long long)(v & 0xFFFFFFFFull) + idx_base);
        if (topk_cost) topk_cost[(size_t)g * topk + i] = (i >= n) ? __int_as_float(0x7fc00000) : key_to_float((uint32_t)(v >> 32));
      } else {
        out_topk[((size_t)g * chunks + chunk) * topk + i] = v;
      }
    }
  }
}

__global__ void gather_best_kernel(const float* __restrict__ a0, const unsigned long long* __restrict__ packed,
                                   int G, int K, long long idx_base, float* __restrict__ best_action,
                                   int* __restrict__ best_idx) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  unsigned long long m = packed[g];
  long long idx = (m == kInvalidPacked) ? 0 : (long long)(m & 0xFFFFFFFFull);
  if (best_idx) best_idx[g] = (int)idx;
  long long local = idx - idx_base;
  if (best_action && local >= 0 && local < K) {
    const float4 a = *reinterpret_cast<const float4*>(a0 + ((size_t)g * K + local) * 4);
    *reinterpret_cast<float4*>(best_action + (size_t)g * 4) = a;
  }
}

// verified[g] = 1 when the refined selection equals the selection from exact costs over ALL K candidates under the
// error model of the single-pass estimate.  The model is an ABSOLUTE bound on the error of the
// terminal state, i.e. an additive bound eta in sqrt-cost space (cost = |s_T - goal|^2 / E, so
// |sqrt(est) - sqrt(exact)| <= |delta s_T| / sqrt(E) whatever the distance to the goal; a relative cost bound would
// be unbounded as s_T approaches the goal).  eta is taken from THIS plan:
//   eta = max( kVerifySafety * max_i |sqrt(est_i) - sqrt(refined_i)|   over the R refined candidates,
//              errc * (rms(goal) + sqrt(est_R))                          a-priori first-order rounding floor (1 sigma),
//              0.5 * eps * sqrt(est_R) )                                 caller's relative floor (eps <= 0: none)
// (the R = 128 observed deviations sample the same near-Gaussian rounding error as the K - R unobserved ones: their
// maximum sits at ~2.7 sigma, the maximum over K = 2^16 .. 2^20 candidates at 4.7 .. 5.3 sigma, so 3 x max covers it with
// ~3 sigma to spare; tests/test_gpu_baseline_cfgs.py::test_verified_flag_is_sound checks eta >= the true maximum.)
// Every candidate outside the list has estimate >= est_R, hence sqrt(exact) >= sqrt(est_R) - eta; if the k-th refined
// cost is strictly below that bound nothing outside can enter the top-k.  One block of 128 threads per goal.
static constexpr float kVerifySafety = 3.0f;
__global__ void __launch_bounds__(128)
verify_refine_kernel(const float* __restrict__ est_topR, const unsigned long long* __restrict__ packed_c,
                     const float* __restrict__ exact_topk, const unsigned long long* __restrict__ packed_min,
                     const float* __restrict__ goal, int E, const float* __restrict__ errc, float eps, int G, int K, int R,
                     int topk, int* __restrict__ verified, float* __restrict__ margin) {
  const int g = blockIdx.x, tid = threadIdx.x;
  if (K <= R) {                                        // the list holds every candidate (or nothing was estimated)
    if (tid == 0) { verified[g] = 1; if (margin) { margin[2 * g] = 0.f; margin[2 * g + 1] = 0.f; } }
    return;
  }
  __shared__ float s_dev[4], s_g2[4];
  __shared__ int s_bad[4];
  float dev = 0.f, g2 = 0.f;
  int bad = 0;
  for (int i = tid; i < R; i += 128) {
    const unsigned long long pc = packed_c[(size_t)g * R + i];
    const float e = est_topR[(size_t)g * R + i];
    if (pc == kPadKey) { bad = 1; continue; }          // K > R: every list slot must hold a candidate
    const float r = key_to_float((uint32_t)(pc >> 32));
    const float d = fabsf(sqrtf(e) - sqrtf(r));
    if (!(d == d) || !(d < 3e38f)) bad = 1;            // NaN / inf anywhere: no proof
    dev = fmaxf(dev, d);
  }
  for (int i = tid; i < E; i += 128) { const float v = __ldg(goal + (size_t)g * E + i); g2 = fmaf(v, v, g2); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dev = fmaxf(dev, __shfl_xor_sync(0xffffffffu, dev, o));
    g2 += __shfl_xor_sync(0xffffffffu, g2, o);
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((tid & 31) == 0) { s_dev[tid >> 5] = dev; s_g2[tid >> 5] = g2; s_bad[tid >> 5] = bad; }
  __syncthreads();
  if (tid == 0) {
    dev = fmaxf(fmaxf(s_dev[0], s_dev[1]), fmaxf(s_dev[2], s_dev[3]));
    g2 = (s_g2[0] + s_g2[1]) + (s_g2[2] + s_g2[3]);
    bad = s_bad[0] | s_bad[1] | s_bad[2] | s_bad[3];
    const float sR = sqrtf(est_topR[(size_t)g * R + R - 1]);
    float eta = kVerifySafety * dev;
    eta = fmaxf(eta, __ldg(errc) * (sqrtf(g2 / (float)E) + sR));
    if (eps > 0.f) eta = fmaxf(eta, 0.5f * eps * sR);
    float kth;
    if (topk > 0) kth = exact_topk[(size_t)g * topk + topk - 1];
    else kth = (packed_min[g] == kInvalidPacked) ? __int_as_float(0x7fc00000) : key_to_float((uint32_t)(packed_min[g] >> 32));
    const float bound = sR - eta;
    verified[g] = (!bad && sqrtf(kth) < bound) ? 1 : 0;     // NaN compares false -> unverified
    if (margin) { margin[2 * g] = bound - sqrtf(kth); margin[2 * g + 1] = eta; }
  }
}

// ------------------------------------------------------------------------------------------------------------
// Single-launch selection (the fast path): one block = one chunk of <= 4096 keys held in REGISTERS (16 per thread,
// coalesced float4 loads issued back to back), block-wide radix select of the topk-th smallest packed key (8-bit
// digits from the top, leading bytes shared by the block's min and max key skipped, histogram updates aggregated per
// warp with match.any), gather of the <= topk survivors and a rank-by-counting sort.  When a goal has several chunks
// the LAST block of the goal (atomic ticket) merges the chunks' survivors the same way: one launch replaces the
// chunk + merge pair.  Bytes: G*K*4 read once, G*(8 + 8*topk) written — the HBM-bound stage of cfg5.
// ------------------------------------------------------------------------------------------------------------
static constexpr int kS2Ept = 16;                       // keys per thread
static constexpr int kS2Chunk = kSelThreads * kS2Ept;   // 4096 keys per block

struct S2Shared {
  unsigned hist[kSelThreads / 32][256];                 // one histogram per warp: plain shared atomics, no cross-warp contention
  unsigned long long list[128];
  unsigned long long wmin[kSelThreads / 32], wmax[kSelThreads / 32];
  unsigned long long prefix, bmin, bmax;
  int need, done, cnt, nreal;
  unsigned ticket;
};

// keys: distinct packed values (pads = kPadKey).  On return sh.list[0..n) holds the n = min(topk, #real) smallest keys
// in ascending order (block-wide; every thread returns n).  *block_min = smallest key (kPadKey if none).
__device__ int block_topk(unsigned long long (&key)[kS2Ept], int topk, S2Shared& sh, unsigned long long* block_min) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long mn = kPadKey, mx = 0ull;
  int nreal = 0;
#pragma unroll
  for (int e = 0; e < kS2Ept; ++e) {
    const bool real = key[e] != kPadKey;
    mn = key[e] < mn ? key[e] : mn;
    mx = (real && key[e] > mx) ? key[e] : mx;
    nreal += real ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
    mn = a < mn ? a : mn;
    mx = b > mx ? b : mx;
    nreal += __shfl_xor_sync(0xffffffffu, nreal, o);
  }
  if (lane == 0) { sh.wmin[warp] = mn; sh.wmax[warp] = mx; sh.hist[0][warp] = (unsigned)nreal; }
  __syncthreads();
  if (tid == 0) {
    unsigned long long a = sh.wmin[0], b = sh.wmax[0];
    int n = (int)sh.hist[0][0];
    for (int w = 1; w < kSelThreads / 32; ++w) {
      a = sh.wmin[w] < a ? sh.wmin[w] : a;
      b = sh.wmax[w] > b ? sh.wmax[w] : b;
      n += (int)sh.hist[0][w];
    }
    sh.bmin = a; sh.bmax = b; sh.nreal = n; sh.cnt = 0;
  }
  __syncthreads();
  *block_min = sh.bmin;
  const int n_out = sh.nreal < topk ? sh.nreal : topk;
  if (n_out == 0) return 0;
  unsigned long long thr;
  if (sh.nreal <= topk) {
    thr = sh.bmax;                                       // everything real is selected
  } else {
    // ---- radix select of the topk-th smallest key: 8-bit digits from the top
    const unsigned long long kmin = sh.bmin, kmax = sh.bmax;
    unsigned long long prefix = 0ull;
    int need = topk, done = 0, shift = 56;
    for (; shift >= 0 && (((kmin ^ kmax) >> shift) == 0ull); shift -= 8) {}   // bytes shared by ALL real keys
    if (shift < 56) prefix = (kmin >> (shift + 8)) << (shift + 8);
    for (; shift >= 0 && !done; shift -= 8) {
#pragma unroll
      for (int w = 0; w < kSelThreads / 32; ++w) sh.hist[w][tid] = 0u;
      __syncthreads();
#pragma unroll
      for (int e = 0; e < kS2Ept; ++e) {
        const unsigned long long kv = key[e];
        const bool act = kv != kPadKey && (shift == 56 || ((kv ^ prefix) >> (shift + 8)) == 0ull);
        if (act) atomicAdd(&sh.hist[warp][(unsigned)(kv >> shift) & 255u], 1u);
      }
      __syncthreads();
      if (warp == 0) {
        unsigned c[8], tot = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          unsigned v = 0;
#pragma unroll
          for (int w = 0; w < kSelThreads / 32; ++w) v += sh.hist[w][lane * 8 + b];
          c[b] = v; tot += v;
        }
        unsigned incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, incl >= (unsigned)need);
        if (lane == __ffs(hit) - 1) {
          unsigned before = incl - tot;
          int b = 0;
          for (; b < 7; ++b) {
            if (before + c[b] >= (unsigned)need) break;
            before += c[b];
          }
          sh.prefix = prefix | ((unsigned long long)(lane * 8 + b) << shift);
          sh.need = need - (int)before;
          sh.done = ((unsigned)(need - (int)before) == c[b]) ? 1 : 0;
        }
      }
      __syncthreads();
      prefix = sh.prefix; need = sh.need; done = sh.done;
      __syncthreads();
    }
    // done at some digit: every key of the chosen bin is selected -> lower bits all ones; else prefix is the exact key
    thr = done ? (prefix | ((shift + 8 > 0) ? ((1ull << (shift + 8)) - 1ull) : 0ull)) : prefix;
  }
#pragma unroll
  for (int e = 0; e < kS2Ept; ++e)
    if (key[e] != kPadKey && key[e] <= thr) {
      const int pos = atomicAdd(&sh.cnt, 1);
      if (pos < 128) sh.list[pos] = key[e];
    }
  __syncthreads();
  const int n = sh.cnt < 128 ? sh.cnt : 128;
  unsigned long long mine = kPadKey;
  int rank = 0;
  if (tid < n) {
    mine = sh.list[tid];
    for (int j = 0; j < n; ++j) rank += (sh.list[j] < mine) ? 1 : 0;
  }
  __syncthreads();
  if (tid < n) sh.list[rank] = mine;
  __syncthreads();
  return n;
}

// MODE 0: fp32 costs [G,K];  MODE 1: packed u64 keys [G,K] (K <= 4096, single chunk)
template <int MODE, bool ARGMIN_ONLY>
__global__ void __launch_bounds__(kSelThreads, ARGMIN_ONLY ? 6 : 4)
select2_kernel(const float* __restrict__ cost, const unsigned long long* __restrict__ packed_in, int K, int topk,
               long long idx_base, unsigned long long* __restrict__ out_min, int* __restrict__ topk_idx,
               float* __restrict__ topk_cost, unsigned long long* __restrict__ part,   // [G][chunks][max(topk,1)]
               unsigned* __restrict__ tickets, const int* __restrict__ skip) {
  if (skip && skip[blockIdx.y]) return;
  __shared__ S2Shared sh;
  const int g = blockIdx.y, chunk = blockIdx.x, chunks = gridDim.x, tid = threadIdx.x;
  const int begin = chunk * kS2Chunk;
  const size_t seg = (size_t)g * (size_t)K;
  unsigned long long key[kS2Ept];
  if (MODE == 0) {
    const float* src = cost + seg;
    float c[kS2Ept];
    const bool vec = ((K & 3) == 0) && ((reinterpret_cast<uintptr_t>(cost) & 15) == 0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i0 = begin + q * 1024 + tid * 4;
      if (vec && i0 + 3 < K) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(src + i0));
        c[4 * q] = v.x; c[4 * q + 1] = v.y; c[4 * q + 2] = v.z; c[4 * q + 3] = v.w;
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) c[4 * q + r] = (i0 + r < K) ? __ldg(src + i0 + r) : 0.f;
      }
    }
    if (ARGMIN_ONLY) {                                   // fold to one running minimum: no key array kept live
      unsigned long long m = kPadKey;
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int i = begin + q * 1024 + tid * 4 + r;
          const unsigned long long kv = i < K ? (((unsigned long long)ordered_key(c[4 * q + r] + 0.0f) << 32) | (unsigned)i) : kPadKey;
          m = kv < m ? kv : m;
        }
#pragma unroll
      for (int e = 0; e < kS2Ept; ++e) key[e] = m;
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int i = begin + q * 1024 + tid * 4 + r;
          key[4 * q + r] = i < K ? (((unsigned long long)ordered_key(c[4 * q + r] + 0.0f) << 32) | (unsigned)i) : kPadKey;
        }
    }
  } else {
#pragma unroll
    for (int e = 0; e < kS2Ept; ++e) {
      const int i = begin + e * kSelThreads + tid;
      key[e] = i < K ? packed_in[seg + i] : kPadKey;
    }
  }
  unsigned long long bmin;
  int n = 0;
  if (ARGMIN_ONLY || topk == 0) {                       // argmin only: warp shuffles + one barrier
    unsigned long long m = key[0];
#pragma unroll
    for (int e = 1; e < kS2Ept; ++e) m = key[e] < m ? key[e] : m;
    m = warp_min_u64(m);
    if ((tid & 31) == 0) sh.wmin[tid >> 5] = m;
    __syncthreads();
    m = sh.wmin[0];
#pragma unroll
    for (int w = 1; w < kSelThreads / 32; ++w) m = sh.wmin[w] < m ? sh.wmin[w] : m;
    bmin = m;
    __syncthreads();
  } else {
    n = block_topk(key, topk, sh, &bmin);
  }
  if (chunks > 1) {
    // publish this chunk's survivors; the last block of the goal merges all chunks
    const int W = topk > 0 ? topk : 1;
    unsigned long long* dst = part + ((size_t)g * chunks + chunk) * W;
    if (topk > 0) { for (int i = tid; i < topk; i += kSelThreads) dst[i] = i < n ? sh.list[i] : kPadKey; }
    else if (tid == 0) dst[0] = bmin;
    __threadfence();
    __syncthreads();
    if (tid == 0) sh.ticket = atomicAdd(tickets + g, 1u);
    __syncthreads();
    if (sh.ticket != (unsigned)(chunks - 1)) return;
    if (tid == 0) tickets[g] = 0u;                       // self-cleaning for the next call
    __threadfence();
    const unsigned long long* src = part + (size_t)g * chunks * W;
    const int total = chunks * W;                        // <= 4096 (host-checked)
#pragma unroll
    for (int e = 0; e < kS2Ept; ++e) {
      const int i = e * kSelThreads + tid;
      key[e] = i < total ? __ldcg(src + i) : kPadKey;
    }
    __syncthreads();
    if (ARGMIN_ONLY) {
      unsigned long long m = key[0];
#pragma unroll
      for (int e = 1; e < kS2Ept; ++e) m = key[e] < m ? key[e] : m;
      m = warp_min_u64(m);
      if ((tid & 31) == 0) sh.wmin[tid >> 5] = m;
      __syncthreads();
      m = sh.wmin[0];
#pragma unroll
      for (int w = 1; w < kSelThreads / 32; ++w) m = sh.wmin[w] < m ? sh.wmin[w] : m;
      bmin = m;
    } else {
      n = block_topk(key, topk, sh, &bmin);
    }
  }
  if (tid == 0) {
    unsigned long long m = bmin;
    // reference rule: only costs strictly below 1e10 may win (mpc_eval.py:122,161-164)
    const float c = key_to_float((uint32_t)(m >> 32));
    if (m == kPadKey || !(c < 1e10f)) m = kInvalidPacked;
    else m = (m & 0xFFFFFFFF00000000ull) | (unsigned long long)((long long)(m & 0xFFFFFFFFull) + idx_base);
    out_min[g] = m;
  }
  for (int i = tid; i < topk; i += kSelThreads) {
    const unsigned long long v = i < n ? sh.list[i] : kPadKey;
    if (topk_idx) topk_idx[(size_t)g * topk + i] = (i >= n) ? -1 : (int)((long long)(v & 0xFFFFFFFFull) + idx_base);
    if (topk_cost) topk_cost[(size_t)g * topk + i] = (i >= n) ? __int_as_float(0x7fc00000) : key_to_float((uint32_t)(v >> 32));
  }
}

// argmin only, many goals of K <= 4096 (BASELINE cfg5: 4096 x 4096 costs = 67 MB): the HBM-bound stage in its purest
// form.  Persistent blocks stride over the goals; the NEXT goal's 16 floats per thread are in flight while the current
// goal is reduced (two register sets), so every SM keeps ~100 KB of loads outstanding instead of paying a block launch
// and a load latency per 16 KB row.
__global__ void __launch_bounds__(kSelThreads, 6)
argmin_rows_kernel(const float* __restrict__ cost, int G, int K, long long idx_base, unsigned long long* __restrict__ out_min,
                   const int* __restrict__ skip) {
  __shared__ unsigned long long s_w[2][kSelThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool vec = ((K & 3) == 0) && ((reinterpret_cast<uintptr_t>(cost) & 15) == 0);
  auto load_row = [&](int g, float (&c)[kS2Ept]) {
    const float* src = cost + (size_t)g * (size_t)K;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int i0 = q * 1024 + tid * 4;
      if (vec && i0 + 3 < K) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(src + i0));      // streamed once: evict first
        c[4 * q] = v.x; c[4 * q + 1] = v.y; c[4 * q + 2] = v.z; c[4 * q + 3] = v.w;
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) c[4 * q + r] = (i0 + r < K) ? __ldcs(src + i0 + r) : __int_as_float(0x7fc00000);
      }
    }
  };
  float cur[kS2Ept], nxt[kS2Ept];
  int g = blockIdx.x;
  if (g < G) load_row(g, cur);
  int it = 0;
  for (; g < G; g += gridDim.x, ++it) {
    const int gn = g + gridDim.x;
    if (gn < G) load_row(gn, nxt);
    // Per thread in the float domain (2-3 instructions per element instead of ~10 for packed 64-bit keys — the packed
    // form made this kernel issue-bound at 79 % issue-active, profiles/r02_ncu_full_argmin_rows_raw.csv): the thread's
    // indices ascend in (q, r), so a strict `<` keeps the first occurrence; NaN and the padded tail never compare
    // below; -0 == +0 tie like their canonical keys.  A thread whose best stays +inf contributes the pad key — any
    // winner >= 1e10 is reported as "none" below anyway (mpc_eval.py:122,161-164), so +inf needs no index.
    float best = __int_as_float(0x7f800000);
    int bi = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float c = cur[4 * q + r];
        const bool lt = c < best;
        best = lt ? c : best;
        bi = lt ? q * 1024 + r : bi;                       // + tid * 4 once, below
      }
    unsigned long long m = best < __int_as_float(0x7f800000)
                               ? (((unsigned long long)ordered_key(best + 0.0f) << 32) | (unsigned)(bi + tid * 4)) : kPadKey;
    m = warp_min_u64(m);
    if (lane == 0) s_w[it & 1][warp] = m;
    __syncthreads();                                     // double-buffered partials: one barrier per goal
    if (tid == 0 && !(skip && skip[g])) {
      m = s_w[it & 1][0];
#pragma unroll
      for (int w = 1; w < kSelThreads / 32; ++w) m = s_w[it & 1][w] < m ? s_w[it & 1][w] : m;
      const float c = key_to_float((uint32_t)(m >> 32));
      if (m == kPadKey || !(c < 1e10f)) m = kInvalidPacked;      // mpc_eval.py:122,161-164
      else m = (m & 0xFFFFFFFF00000000ull) | (unsigned long long)((long long)(m & 0xFFFFFFFFull) + idx_base);
      out_min[g] = m;
    }
#pragma unroll
    for (int e = 0; e < kS2Ept; ++e) cur[e] = nxt[e];
  }
}

// the single-launch path takes: K <= 4096 per goal (one chunk), or chunks * max(topk,1) <= 4096 (in-kernel merge)
static bool select2_eligible(int G, int K, int topk, int* chunks) {
  static const bool off = getenv("NDP_SELECT_OLD") != nullptr;     // diagnostics: force the two-stage kernels
  const int c = (K + kS2Chunk - 1) / kS2Chunk;
  *chunks = c < 1 ? 1 : c;
  if (off || K <= 0) return false;
  return (long long)(*chunks) * (topk > 0 ? topk : 1) <= kS2Chunk && *chunks <= 65535;
}

static void plan_chunks(int G, int K, int* chunks, int* chunk_len) {
  if (K <= 0) { *chunks = 1; *chunk_len = kSelTile; return; }
  int want = 1184 / (G > 0 ? G : 1);
  if (want < 1) want = 1;
  int maxc = (K + kSelTile - 1) / kSelTile;
  int c = want < maxc ? want : maxc;
  if (c < 1) c = 1;
  int len = (K + c - 1) / c;
  len = (len + kSelTile - 1) / kSelTile * kSelTile;
  c = (K + len - 1) / len;
  *chunks = c < 1 ? 1 : c;
  *chunk_len = len;
}

}  // namespace ndp

using namespace ndp;

extern "C" int ndp_select_workspace_bytes(int G, int K, int topk, size_t* bytes) {
  NDP_CHECK_ARG(bytes && G >= 0 && K >= 0 && topk >= 0 && topk <= 128, "select: bad G/K/topk (topk<=128)");
  int chunks, len;
  plan_chunks(G, K, &chunks, &len);
  *bytes = chunks > 1 ? (size_t)G * chunks * (size_t)(topk + 1) * 8 + 256 : 256;
  int c2;
  if (select2_eligible(G, K, topk, &c2) && c2 > 1) {    // tickets [G] + survivors [G][chunks][max(topk,1)]
    const size_t b2 = ((size_t)G * 4 + 255) / 256 * 256 + (size_t)G * c2 * (size_t)(topk > 0 ? topk : 1) * 8 + 256;
    if (b2 > *bytes) *bytes = b2;
  }
  return NDP_OK;
}

static int select_impl(const float* cost, int G, int K, int topk, int64_t idx_base,
                       uint64_t* packed_min, int32_t* topk_idx, float* topk_cost,
                       void* ws, size_t ws_bytes, void* stream, const int* skip) {
  NDP_CHECK_ARG(G >= 0 && K >= 0 && topk >= 0 && topk <= 128, "select: bad G/K/topk (topk<=128)");
  NDP_CHECK_ARG(G == 0 || ((cost || K == 0) && packed_min), "select: null cost/packed_min");
  NDP_CHECK_ARG(idx_base >= 0 && idx_base + (int64_t)K <= 0x7FFFFFFFll, "select: global index must fit int32 (topk_idx / best_idx are int32)");
  NDP_CHECK_ARG(G <= 65535, "select: G must be <= 65535 per call");
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  cudaStream_t s = (cudaStream_t)stream;
  int chunks, len;
  auto* pm = reinterpret_cast<unsigned long long*>(packed_min);
  int c2;
  if (select2_eligible(G, K, topk, &c2)) {              // ONE launch: chunk pass + in-kernel merge by the last block
    unsigned* tickets = nullptr;
    unsigned long long* part = nullptr;
    if (c2 > 1) {
      const size_t toff = ((size_t)G * 4 + 255) / 256 * 256;
      const size_t need2 = toff + (size_t)G * c2 * (size_t)(topk > 0 ? topk : 1) * 8;
      if (ws_bytes < need2 || !ws) { set_error("select: workspace %zu < %zu", ws_bytes, need2); return NDP_ERR_WORKSPACE; }
      NDP_CHECK_ARG(((uintptr_t)ws & 7) == 0, "select: workspace must be 8-byte aligned");
      tickets = reinterpret_cast<unsigned*>(ws);
      part = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(ws) + toff);
      NDP_CUDA(cudaMemsetAsync(tickets, 0, (size_t)G * 4, s));
    }
    if (topk == 0 && c2 == 1 && G >= 64) {
      const int blocks = G < 6 * num_sms() ? G : 6 * num_sms();
      argmin_rows_kernel<<<blocks, kSelThreads, 0, s>>>(cost, G, K, (long long)idx_base, pm, skip);
    } else if (topk == 0)
      select2_kernel<0, true><<<dim3(c2, G), kSelThreads, 0, s>>>(cost, nullptr, K, 0, (long long)idx_base, pm, nullptr, nullptr,
                                                                 part, tickets, skip);
    else
      select2_kernel<0, false><<<dim3(c2, G), kSelThreads, 0, s>>>(cost, nullptr, K, topk, (long long)idx_base, pm, topk_idx,
                                                                  topk_cost, part, tickets, skip);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  plan_chunks(G, K, &chunks, &len);
  if (chunks <= 1) {
    select_kernel<0><<<dim3(1, G), kSelThreads, 0, s>>>(cost, nullptr, K, len > 0 ? len : kSelTile, topk, 1,
                                                        (long long)idx_base, pm, nullptr, topk_idx, topk_cost, skip);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  size_t need = (size_t)G * chunks * (size_t)(topk + 1) * 8;
  if (ws_bytes < need || !ws) { set_error("select: workspace %zu < %zu", ws_bytes, need); return NDP_ERR_WORKSPACE; }
  auto* part_min = reinterpret_cast<unsigned long long*>(ws);
  auto* part_topk = part_min + (size_t)G * chunks;
  select_kernel<0><<<dim3(chunks, G), kSelThreads, 0, s>>>(cost, nullptr, K, len, topk, 0, 0, part_min,
                                                           part_topk, nullptr, nullptr, skip);
  NDP_LAUNCH_CHECK();
  // stage 2: argmin over partial mins and top-k over partial lists (entries keep LOCAL indices)
  if (topk > 0) {
    // merge lists: the min of the merged list equals the min of mins, so one launch does both
    select_kernel<1><<<dim3(1, G), kSelThreads, 0, s>>>(nullptr, part_topk, chunks * topk,
                                                        ((chunks * topk + kSelTile - 1) / kSelTile) * kSelTile,
                                                        topk, 1, (long long)idx_base, pm, nullptr, topk_idx, topk_cost, skip);
  } else {
    select_kernel<1><<<dim3(1, G), kSelThreads, 0, s>>>(nullptr, part_min, chunks,
                                                        kSelTile, 0, 1, (long long)idx_base, pm, nullptr, nullptr, nullptr, skip);
  }
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_select(const float* cost, int G, int K, int topk, int64_t idx_base,
                          uint64_t* packed_min, int32_t* topk_idx, float* topk_cost,
                          void* ws, size_t ws_bytes, void* stream) {
  return select_impl(cost, G, K, topk, idx_base, packed_min, topk_idx, topk_cost, ws, ws_bytes, stream, nullptr);
}

// ---------------------------------------------------------------- ndp_select without any allocation
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the
// workspace of ndp_select lives in a file-scope device array of kSelectStaticBytes that every GPU loading the library
// holds until exit.  In host code the array's NAME is a host-side stand-in, not a device address: its device address is
// resolved once per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels
// get.  Every call on a device shares the array and it keeps the last call's contents: select_impl clears the ticket
// words it needs at zero in stream order before the kernel; calls that overlap on the GPU (different streams) race.
namespace ndp {
constexpr size_t kSelectStaticBytes = (size_t)1 << 20;     // 1 MiB per GPU
static __device__ __align__(256) unsigned char g_select_static_ws[kSelectStaticBytes];

static int select_static_ptr(void** out) {
  static std::mutex mu;
  static void* ptrs[64] = {};
  int dev = 0;
  NDP_CUDA(cudaGetDevice(&dev));
  NDP_CHECK_ARG(dev >= 0 && dev < 64, "select_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_select_static_ws);
    if (e != cudaSuccess || !p) {
      set_error("select_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return NDP_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return NDP_OK;
}
}  // namespace ndp

extern "C" int ndp_select_static_max_bytes(size_t* bytes) {
  NDP_CHECK_ARG(bytes, "select_static_max_bytes: null");
  *bytes = kSelectStaticBytes;
  return NDP_OK;
}

extern "C" int ndp_select_static(const float* cost, int G, int K, int topk, int64_t idx_base, uint64_t* packed_min,
                                 int32_t* topk_idx, float* topk_cost, void* stream) {
  size_t need = 0;
  if (int rc = ndp_select_workspace_bytes(G, K, topk, &need)) return rc;
  if (need > kSelectStaticBytes) {
    set_error("select_static: G=%d K=%d topk=%d needs %zu bytes of scratch, the static array holds %zu; use ndp_select "
              "with a caller workspace", G, K, topk, need, (size_t)kSelectStaticBytes);
    return NDP_ERR_WORKSPACE;
  }
  if (int rc = device_check()) return rc;
  void* ws = nullptr;
  if (int rc = select_static_ptr(&ws)) return rc;
  return select_impl(cost, G, K, topk, idx_base, packed_min, topk_idx, topk_cost, ws, kSelectStaticBytes, stream, nullptr);
}

namespace ndp {
// ndp_select restricted to the goals with skip[g] == 0 (device-side mask, no host decision)
int select_masked(const float* cost, int G, int K, int topk, long long idx_base, unsigned long long* pm, int* topk_idx,
                  float* topk_cost, void* ws, size_t ws_bytes, const int* skip, cudaStream_t s) {
  return select_impl(cost, G, K, topk, idx_base, reinterpret_cast<uint64_t*>(pm), topk_idx, topk_cost, ws, ws_bytes,
                     (void*)s, skip);
}
__global__ void mark_fallback_kernel(int* __restrict__ verified, int G) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < G && verified[g] == 0) verified[g] = 2;      // 2 = recomputed over all K with split operands
}
int mark_fallback(int* verified, int G, cudaStream_t s) {
  mark_fallback_kernel<<<(G + 127) / 128, 128, 0, s>>>(verified, G);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
// selection over packed (okey(cost)<<32 | index) keys, n <= 1024 per goal: one block per goal
int select_packed(const unsigned long long* packed, int G, int n, int topk, long long idx_base,
                  unsigned long long* pm, int* topk_idx, float* topk_cost, cudaStream_t s) {
  NDP_CHECK_ARG(n >= 1 && n <= kSelTile && topk <= 128 && G <= 65535, "select_packed: bad sizes");
  select2_kernel<1, false><<<dim3(1, G), kSelThreads, 0, s>>>(nullptr, packed, n, topk, idx_base, pm, topk_idx, topk_cost,
                                                              nullptr, nullptr, nullptr);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
int verify_refine(const float* est_topR, const unsigned long long* packed_c, const float* exact_topk,
                  const unsigned long long* pm, const float* goal, int E, const float* errc, float eps, int G, int K, int R,
                  int topk, int* verified, float* margin, cudaStream_t s) {
  verify_refine_kernel<<<G, 128, 0, s>>>(est_topR, packed_c, exact_topk, pm, goal, E, errc, eps, G, K, R, topk, verified,
                                         margin);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
}  // namespace ndp

// ---------------------------------------------------------------- sharded plan: record exchange
// One u64 record per goal and rank = everything the other ranks need from a local plan
// (SURVEY 8e): [0] packed argmin key, [1..2] the local winner's action (4 x f32), [3] verified,
// [4..4+topk) packed top-k keys (global indices; missing entries = all ones).  A single
// all-gather of these records replaces min-allreduce + action broadcast + two top-k gathers.
namespace ndp {
__global__ void shard_pack_kernel(const unsigned long long* __restrict__ packed_min, const float* __restrict__ best_action,
                                  const int* __restrict__ verified, const int* __restrict__ topk_idx,
                                  const float* __restrict__ topk_cost, int G, int topk,
                                  unsigned long long* __restrict__ rec) {
  const int W = 4 + topk;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= G * W) return;
  const int g = t / W, w = t % W;
  unsigned long long v;
  if (w == 0) v = packed_min[g];
  else if (w <= 2) {
    const float a = best_action[(size_t)g * 4 + (w - 1) * 2], b = best_action[(size_t)g * 4 + (w - 1) * 2 + 1];
    v = (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32);
  } else if (w == 3) v = verified ? (unsigned long long)(unsigned)verified[g] : 1ull;
  else {
    const int i = topk_idx[(size_t)g * topk + w - 4];
    v = i < 0 ? kPadKey : (((unsigned long long)ordered_key(topk_cost[(size_t)g * topk + w - 4]) << 32) | (unsigned)i);
  }
  rec[t] = v;
}

// rec [P,G,4+topk] -> the global plan.  One block per goal; P*topk <= 1024 keys ranked by counting
// (keys are distinct: indices are global), pad keys rank last in position order.
__global__ void __launch_bounds__(kSelThreads)
shard_merge_kernel(const unsigned long long* __restrict__ rec, int P, int G, int topk,
                   unsigned long long* __restrict__ packed_min, int* __restrict__ best_idx,
                   float* __restrict__ best_action, int* __restrict__ topk_idx, float* __restrict__ topk_cost,
                   int* __restrict__ verified) {
  __shared__ unsigned long long keys[kSelTile];
  const int g = blockIdx.x, W = 4 + topk, n = P * topk;
  for (int i = threadIdx.x; i < n; i += kSelThreads)
    keys[i] = rec[((size_t)(i / topk) * G + g) * W + 4 + i % topk];
  if (threadIdx.x == 0) {
    unsigned long long m = kPadKey;
    int owner = 0, ok = 1;
    for (int r = 0; r < P; ++r) {                        // strict '<': lowest rank == lowest index range wins ties
      const unsigned long long* q = rec + ((size_t)r * G + g) * W;
      if (q[0] < m) { m = q[0]; owner = r; }
      ok &= (q[3] != 0ull) ? 1 : 0;
    }
    if (m == kInvalidPacked) owner = 0;                  // nothing valid anywhere: global index 0 (mpc_eval.py:122,159)
    const unsigned long long* q = rec + ((size_t)owner * G + g) * W;
    packed_min[g] = m;
    if (best_idx) best_idx[g] = (m == kInvalidPacked) ? 0 : (int)(m & 0xFFFFFFFFull);
    if (best_action) {
      best_action[(size_t)g * 4 + 0] = __uint_as_float((unsigned)(q[1] & 0xFFFFFFFFull));
      best_action[(size_t)g * 4 + 1] = __uint_as_float((unsigned)(q[1] >> 32));
      best_action[(size_t)g * 4 + 2] = __uint_as_float((unsigned)(q[2] & 0xFFFFFFFFull));
      best_action[(size_t)g * 4 + 3] = __uint_as_float((unsigned)(q[2] >> 32));
    }
    if (verified) verified[g] = ok;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += kSelThreads) {
    const unsigned long long v = keys[i];
    int r = 0;
    for (int j = 0; j < n; ++j) {
      const unsigned long long u = keys[j];
      r += (u < v) || (u == v && j < i);
    }
    if (r < topk) {
      topk_idx[(size_t)g * topk + r] = (v == kPadKey) ? -1 : (int)(v & 0xFFFFFFFFull);
      topk_cost[(size_t)g * topk + r] = (v == kPadKey) ? __int_as_float(0x7fc00000) : key_to_float((uint32_t)(v >> 32));
    }
  }
}
}  // namespace ndp

extern "C" int ndp_shard_pack(const uint64_t* packed_min, const float* best_action, const int32_t* verified,
                              const int32_t* topk_idx, const float* topk_cost, int G, int topk, uint64_t* rec,
                              void* stream) {
  NDP_CHECK_ARG(G >= 0 && topk >= 0 && topk <= 128, "shard_pack: bad G/topk");
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  NDP_CHECK_ARG(packed_min && best_action && rec && (topk == 0 || (topk_idx && topk_cost)), "shard_pack: null pointer");
  const int n = G * (4 + topk);
  shard_pack_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const unsigned long long*>(packed_min), best_action, verified, topk_idx, topk_cost, G, topk,
      reinterpret_cast<unsigned long long*>(rec));
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

extern "C" int ndp_shard_merge(const uint64_t* rec, int P, int G, int topk, uint64_t* packed_min, int32_t* best_idx,
                               float* best_action, int32_t* topk_idx, float* topk_cost, int32_t* verified,
                               void* stream) {
  NDP_CHECK_ARG(P >= 1 && G >= 0 && topk >= 0 && topk <= 128 && (long long)P * topk <= kSelTile,
                "shard_merge: need P*topk <= 1024 (P=%d topk=%d)", P, topk);
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  NDP_CHECK_ARG(rec && packed_min && (topk == 0 || (topk_idx && topk_cost)), "shard_merge: null pointer");
  shard_merge_kernel<<<G, kSelThreads, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const unsigned long long*>(rec), P, G, topk, reinterpret_cast<unsigned long long*>(packed_min),
      best_idx, best_action, topk_idx, topk_cost, verified);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

// ---------------------------------------------------------------- sharded plan: the exchange as ONE entry point
// pack -> exchange -> merge.  Over the peer window (peer.cuh) it is two launches and no library collective: the pack
// kernel stores each record word straight into EVERY rank's window over NVLink and publishes a flag; the merge kernel
// waits (bounded) for every rank's flag on its own window and merges from local memory.  Records of up to
// kRecSlots plans may be in flight (slot = sequence number % kRecSlots; a PlanRing of depth d keeps at most d + 1
// alive).  Ranks that cannot map each other, or a stream under capture, take pack -> ncclAllGather -> merge.
namespace ndp {
constexpr int kRecSlots = 8;
constexpr int kRecFlag0 = 2;                               // flag slots 2 .. 9 (0, 1: the ndiv phases)
constexpr int kMergeDirect = 256;                          // up to this many keys are merged without shared memory
struct RecWin { size_t off, slot_bytes; };
static RecWin rec_win_layout(size_t win_bytes) {           // second half of the window's capacity, kRecSlots slots
  RecWin r;
  const size_t cap = win_bytes > kWinHeader ? win_bytes - kWinHeader : 0;
  r.slot_bytes = (cap / 2 / kRecSlots) / 256 * 256;
  r.off = kWinHeader + (cap / 2 + 255) / 256 * 256;
  if (r.off + r.slot_bytes * kRecSlots > win_bytes) r.slot_bytes = r.slot_bytes >= 256 ? r.slot_bytes - 256 : 0;
  return r;
}

__global__ void __launch_bounds__(256)
shard_pack_push_kernel(const unsigned long long* __restrict__ packed_min, const float* __restrict__ best_action,
                       const int* __restrict__ verified, const int* __restrict__ topk_idx,
                       const float* __restrict__ topk_cost, int G, int topk, PeerWin w, size_t off, int flag_slot,
                       unsigned epoch) {
  const int W = 4 + topk, n = G * W;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    const int g = t / W, c = t % W;
    unsigned long long v;
    if (c == 0) v = packed_min[g];
    else if (c <= 2) {
      const float a = best_action[(size_t)g * 4 + (c - 1) * 2], b = best_action[(size_t)g * 4 + (c - 1) * 2 + 1];
      v = (unsigned long long)__float_as_uint(a) | ((unsigned long long)__float_as_uint(b) << 32);
    } else if (c == 3) v = verified ? (unsigned long long)(unsigned)verified[g] : 1ull;
    else {
      const int i = topk_idx[(size_t)g * topk + c - 4];
      v = i < 0 ? kPadKey : (((unsigned long long)ordered_key(topk_cost[(size_t)g * topk + c - 4]) << 32) | (unsigned)i);
    }
    for (int r = 0; r < w.world; ++r)                      // rank-major [P][G][W] in every rank's window
      reinterpret_cast<unsigned long long*>(w.peer[r] + off)[(size_t)w.rank * n + t] = v;
  }
  __syncthreads();
  if (threadIdx.x < 32) peer_signal_all(w, flag_slot, epoch);
}

// shard_merge_kernel over the local window after every rank's record has arrived
__global__ void __launch_bounds__(kSelThreads)
shard_wait_merge_kernel(PeerWin w, size_t off, int flag_slot, unsigned epoch, int G, int topk,
                        unsigned long long* __restrict__ packed_min, int* __restrict__ best_idx,
                        float* __restrict__ best_action, int* __restrict__ topk_idx, float* __restrict__ topk_cost,
                        int* __restrict__ verified) {
  // keys staged in (dynamic) shared memory only when there are many (n > kMergeDirect): the bulk rollout leaves
  // 128 B of shared memory per SM, and a merge block without any can run NEXT to a rollout CTA instead of waiting
  // for an SM to drain; n <= 256 keys are re-read from the L2-resident window (n^2 / 256 loads per thread)
  extern __shared__ unsigned long long keys[];
  const int P = w.world;
  if ((int)threadIdx.x < P) peer_wait(w, flag_slot, threadIdx.x, epoch);
  __syncthreads();
  const unsigned long long* rec = reinterpret_cast<const unsigned long long*>(w.peer[w.rank] + off);
  const int g = blockIdx.x, W = 4 + topk, n = P * topk;
  const bool staged = n > kMergeDirect;
  auto key_at = [&](int i) { return ld_peer_u64(rec + ((size_t)(i / topk) * G + g) * W + 4 + i % topk); };
  if (staged)
    for (int i = threadIdx.x; i < n; i += kSelThreads) keys[i] = key_at(i);
  if (threadIdx.x == 0) {
    unsigned long long m = kPadKey;
    int owner = 0, ok = 1;
    for (int r = 0; r < P; ++r) {                        // strict '<': lowest rank == lowest index range wins ties
      const unsigned long long* q = rec + ((size_t)r * G + g) * W;
      const unsigned long long q0 = ld_peer_u64(q);
      if (q0 < m) { m = q0; owner = r; }
      ok &= (ld_peer_u64(q + 3) != 0ull) ? 1 : 0;
    }
    if (m == kInvalidPacked) owner = 0;                  // nothing valid anywhere: global index 0 (mpc_eval.py:122,159)
    const unsigned long long q1 = ld_peer_u64(rec + ((size_t)owner * G + g) * W + 1);
    const unsigned long long q2 = ld_peer_u64(rec + ((size_t)owner * G + g) * W + 2);
    packed_min[g] = m;
    if (best_idx) best_idx[g] = (m == kInvalidPacked) ? 0 : (int)(m & 0xFFFFFFFFull);
    if (best_action) {
      best_action[(size_t)g * 4 + 0] = __uint_as_float((unsigned)(q1 & 0xFFFFFFFFull));
      best_action[(size_t)g * 4 + 1] = __uint_as_float((unsigned)(q1 >> 32));
      best_action[(size_t)g * 4 + 2] = __uint_as_float((unsigned)(q2 & 0xFFFFFFFFull));
      best_action[(size_t)g * 4 + 3] = __uint_as_float((unsigned)(q2 >> 32));
    }
    if (verified) verified[g] = ok;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += kSelThreads) {
    const unsigned long long v = staged ? keys[i] : key_at(i);
    int r = 0;
    for (int j = 0; j < n; ++j) {
      const unsigned long long u = staged ? keys[j] : key_at(j);
      r += (u < v) || (u == v && j < i);
    }
    if (r < topk) {
      topk_idx[(size_t)g * topk + r] = (v == kPadKey) ? -1 : (int)(v & 0xFFFFFFFFull);
      topk_cost[(size_t)g * topk + r] = (v == kPadKey) ? __int_as_float(0x7fc00000) : key_to_float((uint32_t)(v >> 32));
    }
  }
}
}  // namespace ndp

extern "C" int ndp_shard_exchange(ndp_comm* c, const uint64_t* packed_min, const float* best_action,
                                  const int32_t* verified, const int32_t* topk_idx, const float* topk_cost, int G,
                                  int topk, uint64_t* rec, uint64_t* rec_all, uint64_t* out_packed_min,
                                  int32_t* out_best_idx, float* out_best_action, int32_t* out_topk_idx,
                                  float* out_topk_cost, int32_t* out_verified, void* stream) {
  NDP_CHECK_ARG(c && G >= 0 && topk >= 0 && topk <= 128, "shard_exchange: bad communicator/G/topk");
  const int P = ndp_comm_world(c);
  NDP_CHECK_ARG((long long)P * topk <= kSelTile, "shard_exchange: need world*topk <= 1024 (world=%d topk=%d)", P, topk);
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  NDP_CHECK_ARG(packed_min && best_action && out_packed_min && (topk == 0 || (topk_idx && topk_cost && out_topk_idx && out_topk_cost)),
                "shard_exchange: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  const size_t rec_bytes = (size_t)G * (4 + topk) * 8;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  NDP_CUDA(cudaStreamIsCapturing(s, &cap));
  PeerWin pw{};
  if (cap == cudaStreamCaptureStatusNone) {
    // the record region is half of the window's capacity in kRecSlots slots of [world][G][W] words
    if (int rc = comm_window(c, 2 * kRecSlots * ((size_t)P * rec_bytes + 512), s, &pw)) return rc;
  }
  if (pw.world > 0) {
    const RecWin rl = rec_win_layout(c->win_bytes);
    if (rl.slot_bytes < (size_t)P * rec_bytes) { set_error("shard_exchange: window slot %zu < %zu", rl.slot_bytes, (size_t)P * rec_bytes); return NDP_ERR_WORKSPACE; }
    const unsigned epoch = ++c->epoch_rec;
    const int slot = (int)(epoch % kRecSlots);
    const size_t off = rl.off + (size_t)slot * rl.slot_bytes;
    shard_pack_push_kernel<<<1, 256, 0, s>>>(reinterpret_cast<const unsigned long long*>(packed_min), best_action, verified,
                                              topk_idx, topk_cost, G, topk, pw, off, kRecFlag0 + slot, epoch);
    NDP_LAUNCH_CHECK();
    const int nk = P * topk;
    shard_wait_merge_kernel<<<G, kSelThreads, nk > kMergeDirect ? (size_t)nk * 8 : 0, s>>>(pw, off, kRecFlag0 + slot, epoch, G, topk,
                                                      reinterpret_cast<unsigned long long*>(out_packed_min), out_best_idx,
                                                      out_best_action, out_topk_idx, out_topk_cost, out_verified);
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  NDP_CHECK_ARG(rec && rec_all, "shard_exchange: rec / rec_all scratch needed for the NCCL path");
  if (int rc = ndp_shard_pack(packed_min, best_action, verified, topk_idx, topk_cost, G, topk, rec, stream)) return rc;
  if (int rc = ndp_comm_allgather_bytes(c, rec, rec_all, rec_bytes, stream)) return rc;
  return ndp_shard_merge(rec_all, P, G, topk, out_packed_min, out_best_idx, out_best_action, out_topk_idx, out_topk_cost,
                         out_verified, stream);
}

extern "C" int ndp_gather_best_action(const float* a0, const uint64_t* packed_min, int G, int K,
                                      int64_t idx_base, float* best_action, int32_t* best_idx, void* stream) {
  NDP_CHECK_ARG(G >= 0 && packed_min, "gather: bad args");
  if (int rc = device_check()) return rc;
  if (G == 0) return NDP_OK;
  gather_best_kernel<<<(G + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      a0, reinterpret_cast<const unsigned long long*>(packed_min), G, K, (long long)idx_base, best_action, best_idx);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}
