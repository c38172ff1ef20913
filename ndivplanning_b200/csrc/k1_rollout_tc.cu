// K1 (tensor-core modes) — fused generator MLP + T-step latent rollout + terminal cost on
// tcgen05 / TMEM, weights streamed by the TMA engine.  Same contract as k1_rollout_f32.cu
// (reference mpc_eval.py:124-157 + :161, gan.py:79-86, forward_encoder.py:105-114).
//
// One CTA owns 128 candidates (TMEM lanes = candidates) of one goal for the whole horizon.
//   * every layer is D[128 x N] (fp32, TMEM) = A[128 x K] (fp16, TMEM) x W[N x K]^T (fp16, smem);
//     biases, the rank-nd noise update and the rank-4 action update ride in one extra K=16
//     "ext" slab whose A operand lives in shared memory ([noise..,1], [action..,1] or [1,0..]);
//   * epilogue warps turn an accumulator into the next layer's A operand IN PLACE in TMEM
//     (tcgen05.ld -> relu -> fp16 hi[/lo] -> tcgen05.st); hidden activations never touch smem/HBM;
//   * the fp32 state lives in shared memory and is updated with an exact fp32 add per step;
//   * precision: NP=1 single pass fp16; NP=2 activations split hi+lo (2 MMAs per slab);
//     NP=3 activations and weights split (a_hi*w_hi + a_lo*w_hi + a_hi*w_lo): fp32-grade.
//
// TMEM map (512 x 32-bit columns): R0 [0,128) R1 [128,256) DR [256,384) AS [384,512).
// Warp roles: 0-7 epilogue (lane quadrant = warp%4, column half = warp/4), 8 MMA issue, 9 TMA.
#include <stdlib.h>
#include <string.h>
#include "planner.cuh"
#include "tc_common.cuh"

namespace ndp {

using namespace tc;

static constexpr int kTcThreads = 320;
static constexpr uint32_t kColR0 = 0, kColR1 = 128, kColDR = 256, kColAS = 384;
// shared memory map (bytes)
static constexpr uint32_t kOffState = 0;                 // fp32 state, [32 col4][128 rows] float4
static constexpr uint32_t kOffExt = 65536;               // 5 ext slabs x 4 KB: noise hi/lo, action hi/lo, ones
static constexpr uint32_t kExtSlab = 4096;
static constexpr uint32_t kOffRing = kOffExt + 5 * kExtSlab;
static constexpr uint32_t kRingBytes = 122880;           // 6 x 20 KB (NP<=2) or 3 x 40 KB (NP=3)
static constexpr uint32_t kOffItems = kOffRing + kRingBytes;
static constexpr uint32_t kItemsBytes = kTcMaxItems * sizeof(TcItem);
static constexpr uint32_t kOffGoal = kOffItems + kItemsBytes;      // 128 floats
static constexpr uint32_t kOffRed = kOffGoal + 512;                // 128 floats
static constexpr uint32_t kOffBars = kOffRed + 512;                // mbarriers
static constexpr uint32_t kTcSmemBytes = kOffBars + 256;

template <int NP> struct RingCfg { static constexpr uint32_t slot = (NP == 3) ? 40960u : 20480u, nslot = kRingBytes / slot; };

// ------------------------------------------------------------------------------------------
// weight packing: fp32 torch [out,in] -> fp16 hi (/lo) canonical K-major core-matrix layout
//   element (n, kk) of a block with N rows -> half index ((kk/8)*(N/8) + n/8)*64 + (n%8)*8 + kk%8
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void put_split(__half* hi, __half* lo, size_t idx, float v) {
  const __half h = __float2half_rn(v);
  hi[idx] = h;
  if (lo) lo[idx] = __float2half_rn(v - __half2float(h));
}

__global__ void tc_pack_main_kernel(__half* __restrict__ hi, __half* __restrict__ lo, const float* __restrict__ W,
                                    int ld, int n0, int k0, int N, int nslab, int nvalid) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int Kc = nslab * 16;
  if (idx >= N * Kc) return;
  const int n = idx / Kc, kk = idx % Kc;
  const size_t pos = ((size_t)(kk / 8) * (N / 8) + n / 8) * 64 + (n % 8) * 8 + kk % 8;
  put_split(hi, lo, pos, n < nvalid ? W[(size_t)(n0 + n) * ld + k0 + kk] : 0.f);
}

struct ExtCols {
  const float* p[16];   // column k of the ext slab: value for row n is p[k][n * stride[k]]; null -> 0
  int stride[16];
};

__global__ void tc_pack_ext_kernel(__half* __restrict__ hi, __half* __restrict__ lo, int N, int nvalid, ExtCols c) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * 16) return;
  const int n = idx / 16, kk = idx % 16;
  const size_t pos = ((size_t)(kk / 8) * (N / 8) + n / 8) * 64 + (n % 8) * 8 + kk % 8;
  const float v = (c.p[kk] && n < nvalid) ? c.p[kk][(size_t)n * c.stride[kk]] : 0.f;
  put_split(hi, lo, pos, v);
}

// per-goal ext block of generator fc1: columns [w1 noise cols (nd) | b1 + W1g . goal | 0 ...], N = 128.
// The goal half of fc1 is constant per plan (the reference recomputes it per candidate and step,
// mpc_eval.py:133,136-138).  One block per goal, one warp per output row.
__global__ void tc_goal_ext_kernel(const float* __restrict__ W1, const float* __restrict__ b1,
                                   const float* __restrict__ goal, int E, int nd, int with_lo,
                                   __half* __restrict__ gext, size_t stride_halfs,
                                   unsigned* __restrict__ clear_flags, int n_flags) {
  const int g = blockIdx.x;
  const int lane = threadIdx.x & 31;
  if (blockIdx.x == 0 && blockIdx.y == 0)                  // hand-over flags of the split pairs (two-tile kernel)
    for (int i = threadIdx.x; i < n_flags; i += blockDim.x) clear_flags[i] = 0u;
  const int C1 = 2 * E + nd;
  __half* hi = gext + (size_t)g * stride_halfs;
  __half* lo = with_lo ? hi + 128 * 16 : nullptr;
  {
    const int n = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5);   // one warp per output row
    float acc = 0.f;
    for (int e = lane; e < E; e += 32)
      acc = fmaf(__ldg(W1 + (size_t)n * C1 + E + e), __ldg(goal + (size_t)g * E + e), acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    const float bias = acc + __ldg(b1 + n);
    if (lane < 16) {
      const int kk = lane;
      float v = 0.f;
      if (kk < nd) v = __ldg(W1 + (size_t)n * C1 + 2 * E + kk);
      else if (kk == nd) v = bias;
      const size_t pos = ((size_t)(kk / 8) * 16 + n / 8) * 64 + (n % 8) * 8 + kk % 8;
      put_split(hi, lo, pos, v);
    }
  }
}

// ------------------------------------------------------------------------------------------
// epilogue helpers (thread = one candidate row; `lane_addr` has the TMEM lane field set)
// ------------------------------------------------------------------------------------------
// 64 accumulator columns -> relu -> fp16 hi[/lo] written in place: hi at [0,32), lo at [32,64)
template <int NP>
__device__ __forceinline__ void relu_inplace64(uint32_t addr) {
  uint32_t v[32], w[32];
  tmem_ld32(addr, v);
  tmem_ld32(addr + 32, w);
  tmem_ld_wait();
  uint32_t hi[32];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float x0 = fmaxf(__uint_as_float(v[2 * j]), 0.f), x1 = fmaxf(__uint_as_float(v[2 * j + 1]), 0.f);
    hi[j] = pack_hi(x0, x1);
    if (NP >= 2) v[j] = pack_lo(x0, x1, hi[j]);   // v[j] (j<16) is dead once pair j (<= 2j) was consumed
  }
  // second half; lo words go to w[] the same way
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float x0 = fmaxf(__uint_as_float(w[2 * j]), 0.f), x1 = fmaxf(__uint_as_float(w[2 * j + 1]), 0.f);
    hi[16 + j] = pack_hi(x0, x1);
    if (NP >= 2) w[j] = pack_lo(x0, x1, hi[16 + j]);
  }
  tmem_st16(addr, hi);
  tmem_st16(addr + 16, hi + 16);
  if (NP >= 2) {
    tmem_st16(addr + 32, v);
    tmem_st16(addr + 48, w);
  }
}
// 32 accumulator columns (N=64 layer): hi at [0,16), lo at [16,32)
template <int NP>
__device__ __forceinline__ void relu_inplace32(uint32_t addr) {
  uint32_t v[32];
  tmem_ld32(addr, v);
  tmem_ld_wait();
  uint32_t hi[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float x0 = fmaxf(__uint_as_float(v[2 * j]), 0.f), x1 = fmaxf(__uint_as_float(v[2 * j + 1]), 0.f);
    hi[j] = pack_hi(x0, x1);
    if (NP >= 2) v[j] = pack_lo(x0, x1, hi[j]);
  }
  tmem_st16(addr, hi);
  if (NP >= 2) tmem_st16(addr + 16, v);
}

// state row half (64 fp32 values, already in registers as s[]) -> AS operand region (hi [0,32), lo [32,64))
template <int NP>
__device__ __forceinline__ void store_state_operand(uint32_t addr, const float (&s)[64]) {
  uint32_t hi[32], lo[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    hi[j] = pack_hi(s[2 * j], s[2 * j + 1]);
    if (NP >= 2) lo[j] = pack_lo(s[2 * j], s[2 * j + 1], hi[j]);
  }
  tmem_st16(addr, hi);
  tmem_st16(addr + 16, hi + 16);
  if (NP >= 2) {
    tmem_st16(addr + 32, lo);
    tmem_st16(addr + 48, lo + 16);
  }
}

// one row of an ext slab in shared memory: 16 values -> fp16 hi (and lo) in the canonical layout
__device__ __forceinline__ void store_ext_row(uint8_t* slab_hi, uint8_t* slab_lo, int m, const float (&x)[16], bool second_chunk) {
  const uint32_t off = (uint32_t)(m >> 3) * 128 + (uint32_t)(m & 7) * 16;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (c == 1 && !second_chunk) break;
    uint4 h, l;
    h.x = pack_hi(x[8 * c + 0], x[8 * c + 1]); l.x = pack_lo(x[8 * c + 0], x[8 * c + 1], h.x);
    h.y = pack_hi(x[8 * c + 2], x[8 * c + 3]); l.y = pack_lo(x[8 * c + 2], x[8 * c + 3], h.y);
    h.z = pack_hi(x[8 * c + 4], x[8 * c + 5]); l.z = pack_lo(x[8 * c + 4], x[8 * c + 5], h.z);
    h.w = pack_hi(x[8 * c + 6], x[8 * c + 7]); l.w = pack_lo(x[8 * c + 6], x[8 * c + 7], h.w);
    *reinterpret_cast<uint4*>(slab_hi + c * 2048 + off) = h;
    *reinterpret_cast<uint4*>(slab_lo + c * 2048 + off) = l;
  }
}

// ------------------------------------------------------------------------------------------
// MMA issue of one stream item (one ring slot): NMAIN main slabs (A in TMEM) + optional ext slab
// (A in smem).  Everything but the slot position is a compile-time constant so the issuing lane
// spends ~10 instructions per MMA; the item ORDER here must match the host builder's item table.
// ------------------------------------------------------------------------------------------
#ifdef NDP_TC_PROF
__device__ unsigned long long g_tc_prof[8];
__device__ unsigned long long g_tc_item_wait[64], g_tc_a_wait[16];   // [0] wait_a [1] wait_full [2] issue [3] total (MMA warp); [4] epi wait_d [5] epi work [6] epi total
#define PROF_T0(v) long long v = clock64()
#define PROF_ADD(slot, v) do { long long n__ = clock64(); st_prof[slot] += n__ - v; v = n__; } while (0)
#else
#define PROF_T0(v)
#define PROF_ADD(slot, v)
#endif

struct IssueState {
  uint32_t slot, par;        // ring position / parity of the "full" barriers
  uint32_t bars;             // smem address of the barrier block
  uint32_t ring16;           // ring base >> 4
  uint32_t ext16[5];         // (ext slab address >> 4) | LBO bits: noise hi, noise lo, action hi, action lo, ones
  uint32_t ds;               // "D full" events signalled so far
  uint32_t next_ready;       // result of the early try_wait on the current slot (issued one item ahead)
#ifdef NDP_TC_PROF
  long long prof[4];
  int item, aev;
#endif
};

template <int NP, int N, int NMAIN, int EXT, bool SIGNAL_D, int SPC = 4, int CPH = 64>
__device__ __forceinline__ void tc_issue(IssueState& st, uint32_t d, uint32_t a, uint32_t lo_delta, uint32_t first) {
  constexpr uint32_t kSlot = RingCfg<NP>::slot, kNSlot = RingCfg<NP>::nslot;
  constexpr uint32_t kDescHi = 0x4000u | (128u >> 4);          // descriptor version 1, SBO = 128 B
  constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  constexpr uint32_t step = 2 * N;                              // one slab = N*32 bytes, in 16-byte units
  constexpr uint32_t lo_w = NMAIN * step;                       // hi -> lo weight block distance (NP=3)
#ifdef NDP_TC_PROF
  long long* st_prof = st.prof;
#endif
  PROF_T0(tp);
  if (!st.next_ready) mbar_wait(st.bars + 8 * st.slot, st.par);
#ifdef NDP_TC_PROF
  if ((threadIdx.x & 31) == 0) atomicAdd(&g_tc_item_wait[st.item & 63], (unsigned long long)(clock64() - tp));
  ++st.item;
#endif
  PROF_ADD(1, tp);
  fence_after_sync();
  if (elect_one()) {
    uint32_t blo = (st.ring16 + st.slot * (kSlot >> 4)) | ((uint32_t)N << 16);   // LBO = N*16 bytes
    uint32_t acc = first ? 0u : 1u;
#pragma unroll
    for (int s = 0; s < NMAIN; ++s) {
      const uint64_t bd = pack64(blo + s * step, kDescHi);
      const uint32_t as = a + (s / SPC) * CPH + (s % SPC) * 8;   // operand halves are CPH columns apart
      mma_ts(d, as, bd, kIdesc, acc);
      acc = 1;
      if (NP >= 2) mma_ts(d, as + lo_delta, bd, kIdesc, 1);
      if (NP == 3) mma_ts(d, as, pack64(blo + s * step + lo_w, kDescHi), kIdesc, 1);
    }
    if (EXT) {
      const uint32_t eb = blo + ((NP == 3) ? 2 * lo_w : lo_w);  // ext block follows [main hi][main lo]
      const uint64_t bd = pack64(eb, kDescHi);
      const uint64_t ad = pack64(st.ext16[EXT == 1 ? 0 : (EXT == 2 ? 2 : 4)], kDescHi);
      mma_ss(d, ad, bd, kIdesc, acc);
      if (NP >= 2 && EXT != 3) mma_ss(d, pack64(st.ext16[EXT == 1 ? 1 : 3], kDescHi), bd, kIdesc, 1);
      if (NP == 3) mma_ss(d, ad, pack64(eb + step, kDescHi), kIdesc, 1);
    }
    mma_commit(st.bars + 8 * (6 + st.slot));
    if (SIGNAL_D) mma_commit(st.bars + 8 * (12 + (st.ds & 1)));
  }
  __syncwarp();
  if (SIGNAL_D) ++st.ds;
  if (++st.slot == kNSlot) { st.slot = 0; st.par ^= 1; }
  // probe the next slot now: the ~70-cycle try_wait latency overlaps the bookkeeping / next wait_a
  st.next_ready = mbar_try_wait(st.bars + 8 * st.slot, st.par) ? 1u : 0u;
  PROF_ADD(2, tp);
}

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <int NP>
__global__ void __launch_bounds__(kTcThreads, 1) rollout_tc_kernel(const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t kSlot = RingCfg<NP>::slot, kNSlot = RingCfg<NP>::nslot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = blockIdx.x / p.tiles_per_goal;
  const int m0 = (blockIdx.x % p.tiles_per_goal) * 128;
  if (p.skip && __ldg(p.skip + g)) return;             // CTA-uniform, before any barrier / TMEM allocation

  float4* S4 = reinterpret_cast<float4*>(smem + kOffState);
  uint8_t* ext_noise_hi = smem + kOffExt;
  uint8_t* ext_noise_lo = ext_noise_hi + kExtSlab;
  uint8_t* ext_act_hi = ext_noise_lo + kExtSlab;
  uint8_t* ext_act_lo = ext_act_hi + kExtSlab;
  uint8_t* ext_ones = ext_act_lo + kExtSlab;
  TcItem* items = reinterpret_cast<TcItem*>(smem + kOffItems);
  float* s_goal = reinterpret_cast<float*>(smem + kOffGoal);
  float* s_red = reinterpret_cast<float*>(smem + kOffRed);
  const uint32_t bars = smem_u32(smem + kOffBars);
  // barrier map (8 bytes each): full[0..5] empty[6..11] d_full[12,13] a_ready[14,15]; tmem base at +128
  auto bar_full = [&](uint32_t s) { return bars + 8 * s; };
  auto bar_empty = [&](uint32_t s) { return bars + 8 * (6 + s); };
  auto bar_dfull = [&](uint32_t e) { return bars + 8 * (12 + (e & 1)); };
  auto bar_aready = [&](uint32_t e) { return bars + 8 * (14 + (e & 1)); };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffBars + 128);

  // ---- one-time setup ----------------------------------------------------------------------
  for (int i = tid; i < p.n_items * 8; i += kTcThreads)
    reinterpret_cast<uint32_t*>(items)[i] = reinterpret_cast<const uint32_t*>(p.items)[i];
  for (int i = tid; i < 5 * (int)kExtSlab / 16; i += kTcThreads)
    reinterpret_cast<uint4*>(smem + kOffExt)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 128) s_goal[tid] = __ldg(p.goal + (size_t)g * 128 + tid);
  if (tid == 0) {
    for (uint32_t s = 0; s < kNSlot; ++s) { mbar_init(bar_full(s), 1); mbar_init(bar_empty(s), 1); }
    mbar_init(bar_dfull(0), 1); mbar_init(bar_dfull(1), 1);
    mbar_init(bar_aready(0), 8); mbar_init(bar_aready(1), 8);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc512(smem_u32((const void*)tmem_slot));
  __syncthreads();
  // "ones" slab: element k=0 of every row is 1.0 (fp16 0x3C00); rows are 16 B apart inside a core matrix
  if (tid < 128) *reinterpret_cast<uint16_t*>(ext_ones + (tid >> 3) * 128 + (tid & 7) * 16) = 0x3C00;
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp == 9) {
    // ===================== TMA producer: stream weight blocks into the ring =====================
    if (lane == 0) {
      uint32_t it = 0;
      const uint8_t* gsrc = p.gext + (size_t)g * p.gext_stride;
      for (int t = 0; t < p.T; ++t) {
        for (int i = 0; i < p.n_items; ++i, ++it) {
          const uint32_t slot = it % kNSlot, par = (it / kNSlot) & 1;
          const uint32_t bytes = items[i].bytes, eb = items[i].ext_bytes;
          mbar_wait(bar_empty(slot), par ^ 1);
          mbar_expect_tx(bar_full(slot), bytes + eb);
          const uint32_t dst = smem_u32(smem + kOffRing + slot * kSlot);
          bulk_g2s(dst, p.stream + items[i].src_off, bytes, bar_full(slot));
          if (eb) bulk_g2s(dst + bytes, gsrc, eb, bar_full(slot));
        }
      }
    }
  } else if (warp == 8) {
    // ===================== MMA issuer (whole warp converged; one elected lane issues) ===========
    IssueState st;
    st.slot = 0; st.par = 0; st.next_ready = 0; st.bars = bars; st.ring16 = smem_u32(smem + kOffRing) >> 4; st.ds = 0;
    const uint32_t ext_lbo = (2048u >> 4) << 16;               // A ext slabs: LBO = 2 KB, SBO = 128 B
    st.ext16[0] = (smem_u32(ext_noise_hi) >> 4) | ext_lbo; st.ext16[1] = (smem_u32(ext_noise_lo) >> 4) | ext_lbo;
    st.ext16[2] = (smem_u32(ext_act_hi) >> 4) | ext_lbo;   st.ext16[3] = (smem_u32(ext_act_lo) >> 4) | ext_lbo;
    st.ext16[4] = (smem_u32(ext_ones) >> 4) | ext_lbo;
    uint32_t aw = 0;
#ifdef NDP_TC_PROF
    st.prof[0] = st.prof[1] = st.prof[2] = st.prof[3] = 0;
    long long* st_prof = st.prof;
    const long long t_begin = clock64();
#endif
    auto wait_a = [&]() {
      PROF_T0(tw);
      mbar_wait(bar_aready(aw), (aw >> 1) & 1);
      ++aw;
      fence_after_sync();
#ifdef NDP_TC_PROF
      if (lane == 0) atomicAdd(&g_tc_a_wait[st.aev & 15], (unsigned long long)(clock64() - tw));
      ++st.aev;
#endif
      PROF_ADD(0, tw);
    };
    const uint32_t R0 = tmem + kColR0, R1 = tmem + kColR1, DR = tmem + kColDR, AS = tmem + kColAS;
    const int nch = p.n_hid_chunks;
    for (int t = 0; t < p.T; ++t) {
#ifdef NDP_TC_PROF
      st.item = 0; st.aev = 0;
#endif
      // generator fc1 (A = state operand, ext = [noise | per-goal bias])
      wait_a();
      tc_issue<NP, 128, 4, 0, false>(st, R0, AS, 32, 1);
      tc_issue<NP, 128, 4, 1, true>(st, R0, AS + 64, 32, 0);
      // fc2 (N=64)
      wait_a();
      tc_issue<NP, 64, 8, 3, true>(st, R1, R0, 32, 1);
      // fc3 (K=64: operand halves are 32 columns wide, hi 16 + lo 16)
      wait_a();
      tc_issue<NP, 128, 4, 3, true, 2, 32>(st, R0, R1, 16, 1);
      // fc4 as two N=128 halves
      wait_a();
      tc_issue<NP, 128, 4, 0, false>(st, R1, R0, 32, 1);
      tc_issue<NP, 128, 4, 3, true>(st, R1, R0 + 64, 32, 0);
      tc_issue<NP, 128, 4, 0, false>(st, DR, R0, 32, 1);
      tc_issue<NP, 128, 4, 3, true>(st, DR, R0 + 64, 32, 0);
      // fc5 (N=16, K=256 over R1|DR)
      wait_a();
      tc_issue<NP, 16, 16, 3, true>(st, R0, R1, 32, 1);
      // forward model: F1[0], then F1[c] / F2[c-1] interleaved, then F2[last]
      wait_a();
      tc_issue<NP, 128, 4, 0, false>(st, R0, AS, 32, 1);
      tc_issue<NP, 128, 4, 2, true>(st, R0, AS + 64, 32, 0);
      for (int c = 1; c < nch; ++c) {
        const uint32_t Rc = (c & 1) ? R1 : R0, Rp = (c & 1) ? R0 : R1;
        tc_issue<NP, 128, 4, 0, false>(st, Rc, AS, 32, 1);
        tc_issue<NP, 128, 4, 2, true>(st, Rc, AS + 64, 32, 0);
        wait_a();
        tc_issue<NP, 128, 4, 0, false>(st, DR, Rp, 32, c == 1);
        tc_issue<NP, 128, 4, 0, false>(st, DR, Rp + 64, 32, 0);
      }
      {
        const uint32_t Rp = ((nch - 1) & 1) ? R1 : R0;
        wait_a();
        tc_issue<NP, 128, 4, 0, false>(st, DR, Rp, 32, nch == 1);
        tc_issue<NP, 128, 4, 3, true>(st, DR, Rp + 64, 32, 0);
      }
    }
#ifdef NDP_TC_PROF
    if (lane == 0) {
      atomicAdd(&g_tc_prof[0], (unsigned long long)st.prof[0]);
      atomicAdd(&g_tc_prof[1], (unsigned long long)st.prof[1]);
      atomicAdd(&g_tc_prof[2], (unsigned long long)st.prof[2]);
      atomicAdd(&g_tc_prof[3], (unsigned long long)(clock64() - t_begin));
    }
#endif
  } else {
    // ===================== epilogue warps 0..7 =====================================================
    const int q = warp & 3, h = warp >> 2;
    const int m = q * 32 + lane;                       // candidate row = TMEM lane
    int cand = m0 + m;
    bool valid = cand < p.K;
    if (p.list) {                                         // refine pass: explicit candidate list
      cand = (m0 + m < p.R) ? __ldg(p.list + (size_t)g * p.R + m0 + m) : -1;
      valid = cand >= 0 && cand < p.K;
    }
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    uint32_t dw = 0, ae = 0;
#ifdef NDP_TC_PROF
    long long st_prof[4] = {0, 0, 0, 0};
    const long long t_begin = clock64();
#endif
    auto wait_d = [&]() {
      PROF_T0(tw);
      mbar_wait(bar_dfull(dw), (dw >> 1) & 1);
      ++dw;
      fence_after_sync();
      PROF_ADD(0, tw);
    };
    auto signal_a = [&](bool wrote_smem = false) {
      tmem_st_wait();
      if (wrote_smem) fence_proxy_async();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_aready(ae));
      ++ae;
    };
    auto load_noise = [&](int t, float (&x)[16]) {
#pragma unroll
      for (int k = 0; k < 16; ++k) x[k] = 0.f;
      if (h == 0 && t < p.T) {
        const float* src = p.noise + (((size_t)t * p.G + g) * p.K + cand) * p.nd;
#pragma unroll
        for (int k = 0; k < 15; ++k)
          if (k < p.nd && valid) x[k] = __ldg(src + k);
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (k == p.nd) x[k] = 1.f;
      }
    };
    // ---- init event: state <- e0, AS <- split(e0), noise ext of step 0
    {
      float s[64];
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(p.e0 + (size_t)g * 128) + h * 16 + c);
        S4[(h * 16 + c) * 128 + m] = v;
        s[4 * c] = v.x; s[4 * c + 1] = v.y; s[4 * c + 2] = v.z; s[4 * c + 3] = v.w;
      }
      store_state_operand<NP>(lane_base + kColAS + 64 * h, s);
      float nz[16];
      load_noise(0, nz);
      if (h == 0) store_ext_row(ext_noise_hi, ext_noise_lo, m, nz, p.nd >= 8);
      signal_a(true);
    }
    float nz[16];
    for (int t = 0; t < p.T; ++t) {
      load_noise(t + 1, nz);                                      // prefetch next step's noise
      // ---- G1 (N=128, R0), G2 (N=64, R1), G3 (N=128, R0): relu -> next A operand in place
      wait_d(); relu_inplace64<NP>(lane_base + kColR0 + 64 * h); signal_a();
      wait_d(); relu_inplace32<NP>(lane_base + kColR1 + 32 * h); signal_a();
      wait_d(); relu_inplace64<NP>(lane_base + kColR0 + 64 * h); signal_a();
      // ---- G4 (two N=128 halves in R1 and DR)
      wait_d();
      relu_inplace64<NP>(lane_base + kColR1 + 64 * h);
      wait_d();
      relu_inplace64<NP>(lane_base + kColDR + 64 * h);
      signal_a();
      // ---- G5 (N=16 in R0, 4 real columns): action -> ext slab; step-0 actions are the plan's a0
      wait_d();
      if (h == 0) {
        uint32_t av[4];
        tmem_ld4(lane_base + kColR0, av);
        tmem_ld_wait();
        float x[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) x[k] = 0.f;
        x[0] = __uint_as_float(av[0]); x[1] = __uint_as_float(av[1]);
        x[2] = __uint_as_float(av[2]); x[3] = __uint_as_float(av[3]);
        x[4] = 1.f;
        store_ext_row(ext_act_hi, ext_act_lo, m, x, false);
        if (t == 0 && valid)   // mpc_eval.py:150-151
          *reinterpret_cast<float4*>(p.a0 + ((size_t)g * p.K + cand) * 4) = make_float4(x[0], x[1], x[2], x[3]);
      }
      signal_a(true);
      // ---- forward model hidden chunks: F1[c] accumulator (R0/R1 alternating) -> relu -> A of F2[c]
      for (int c = 0; c < p.n_hid_chunks; ++c) {
        wait_d();
        relu_inplace64<NP>(lane_base + ((c & 1) ? kColR1 : kColR0) + 64 * h);
        signal_a();
      }
      // ---- END: state += residual (fp32, exact add; forward_encoder.py:113-114), new A operand, next noise
      float4 sv[16];
#pragma unroll
      for (int c = 0; c < 16; ++c) sv[c] = S4[(h * 16 + c) * 128 + m];     // state loads overlap the wait
      wait_d();
      {
        uint32_t r0[32], r1[32];
        tmem_ld32(lane_base + kColDR + 64 * h, r0);
        tmem_ld32(lane_base + kColDR + 64 * h + 32, r1);
        tmem_ld_wait();
        float s[64];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          float4 v = sv[c];
          const uint32_t* r = (c < 8) ? (r0 + 4 * c) : (r1 + 4 * (c - 8));
          v.x += __uint_as_float(r[0]); v.y += __uint_as_float(r[1]);
          v.z += __uint_as_float(r[2]); v.w += __uint_as_float(r[3]);
          s[4 * c] = v.x; s[4 * c + 1] = v.y; s[4 * c + 2] = v.z; s[4 * c + 3] = v.w;
        }
        if (t + 1 < p.T) {
          store_state_operand<NP>(lane_base + kColAS + 64 * h, s);
          if (h == 0) store_ext_row(ext_noise_hi, ext_noise_lo, m, nz, p.nd >= 8);
          signal_a(true);
          // the fp32 state goes back to shared memory off the critical path (only this thread reads it)
#pragma unroll
          for (int c = 0; c < 16; ++c)
            S4[(h * 16 + c) * 128 + m] = make_float4(s[4 * c], s[4 * c + 1], s[4 * c + 2], s[4 * c + 3]);
        } else {
          // terminal cost: mean over E of squared error (nn.MSELoss, mpc_eval.py:161)
          float cs = 0.f;
#pragma unroll
          for (int c = 0; c < 64; ++c) {
            const float df = s[c] - s_goal[64 * h + c];
            cs = fmaf(df, df, cs);
          }
          if (h == 1) s_red[m] = cs;
          asm volatile("bar.sync 1, 256;" ::: "memory");
          if (h == 0) {
            const float cv = (cs + s_red[m]) * (1.0f / 128.0f);
            if (valid) p.cost[(size_t)g * p.K + cand] = cv;
            if (p.packed_c && m0 + m < p.R)
              p.packed_c[(size_t)g * p.R + m0 + m] =
                  valid ? (((unsigned long long)ordered_key(cv + 0.0f) << 32) | (unsigned)cand) : ~0ull;
          }
          if (p.state_T && valid) {
            float4* o = reinterpret_cast<float4*>(p.state_T + ((size_t)g * p.K + cand) * 128) + h * 16;
#pragma unroll
            for (int c = 0; c < 16; ++c) o[c] = make_float4(s[4 * c], s[4 * c + 1], s[4 * c + 2], s[4 * c + 3]);
          }
        }
      }
    }
  #ifdef NDP_TC_PROF
    if (warp == 0 && lane == 0) {
      atomicAdd(&g_tc_prof[4], (unsigned long long)st_prof[0]);
      atomicAdd(&g_tc_prof[6], (unsigned long long)(clock64() - t_begin));
    }
#endif
  }
#ifdef NDP_TC_PROF
  // (epilogue counters are accumulated by warp 0 lane 0 only: representative)
#endif
  fence_before_sync();
  __syncthreads();
  if (warp == 8) tmem_dealloc512(tmem);
}

// ------------------------------------------------------------------------------------------
// host side: weight stream + item table of one precision mode
// ------------------------------------------------------------------------------------------
struct TcPlan {
  int np, nd, H, n_items, n_chunks;
  uint8_t* stream;        // device
  size_t stream_bytes;
  TcItem* items;          // device
  const float* W1;        // device, torch layout (planner-owned copies)
  const float* b1;
  uint32_t gext_stride;   // bytes per goal
  // two-tile bulk kernel (np == 1): its own item order / chunking (rollout_tc2_kernel)
  uint8_t* stream2;
  TcItem* items2;
  int n_items2, n_chunks2;
};

namespace {

struct Builder {
  int np;
  cudaStream_t s;
  uint8_t* stream;
  size_t off = 0;
  TcItem items[kTcMaxItems];
  int n = 0;
  int rc = NDP_OK;

  // one item: `nslab` main slabs [j0, j0+nslab) of W rows [n0, n0+N) (rows >= nvalid are zero), optional ext slab.
  // `a_reg`/`cph`: TMEM region of the A operand and the D columns per epilogue half that produced it.
  void add(const float* W, int ld, int n0, int nvalid, int kcol0, int N, int j0, int nslab, int d_col, int a_reg,
           int cph, bool first, bool wait_a, bool signal_d, int ext_kind, const ExtCols* ext) {
    if (rc) return;
    if (n >= kTcMaxItems) { set_error("tensor-core plan: too many stream items (H too large)"); rc = NDP_ERR_ARG; return; }
    const int spc = cph / 16;                                   // slabs per epilogue half
    const size_t slab = (size_t)N * 32, main_hi = slab * nslab;
    TcItem& I = items[n++];
    memset(&I, 0, sizeof(I));
    I.src_off = (uint32_t)off;
    I.idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    I.N = (uint16_t)N; I.d_col = (uint16_t)d_col;
    I.a_col = (uint16_t)(a_reg + (j0 / spc) * cph + (j0 % spc) * 8);
    I.lo_delta = (uint16_t)(cph / 2);
    I.n_main = (uint8_t)nslab; I.ext_kind = (uint8_t)ext_kind; I.first = first; I.wait_a = wait_a; I.signal_d = signal_d;
    __half* hi = reinterpret_cast<__half*>(stream + off);
    __half* lo = (np == 3) ? reinterpret_cast<__half*>(stream + off + main_hi) : nullptr;
    const int total = N * nslab * 16;
    tc_pack_main_kernel<<<(total + 255) / 256, 256, 0, s>>>(hi, lo, W, ld, n0, kcol0 + j0 * 16, N, nslab, nvalid);
    count_launch();
    size_t bytes = main_hi * (np == 3 ? 2 : 1);
    if (ext_kind == 1) {
      I.ext_bytes = (uint32_t)(slab * (np == 3 ? 2 : 1));       // per-goal block, built per call
    } else if (ext_kind) {
      __half* ehi = reinterpret_cast<__half*>(stream + off + bytes);
      __half* elo = (np == 3) ? reinterpret_cast<__half*>(stream + off + bytes + slab) : nullptr;
      tc_pack_ext_kernel<<<(N * 16 + 255) / 256, 256, 0, s>>>(ehi, elo, N, nvalid, *ext);
      count_launch();
      bytes += slab * (np == 3 ? 2 : 1);
    }
    I.bytes = (uint32_t)bytes;
    off += (bytes + 127) / 128 * 128;
    if (cudaGetLastError() != cudaSuccess) { set_error("tensor-core plan: pack kernel launch failed"); rc = NDP_ERR_CUDA; }
  }
};

ExtCols bias_only(const float* b) {
  ExtCols c;
  memset(&c, 0, sizeof(c));
  c.p[0] = b; c.stride[0] = 1;
  return c;
}

}  // namespace

static int tc2_build(TcPlan* P, int E, int nd, int H, const float* const* Wt, const float* const* bt, cudaStream_t s);

int tc_plan_create(TcPlan** out, int np, int E, int nd, int H, const float* const* Wt, const float* const* bt,
                   cudaStream_t s) {
  NDP_CHECK_ARG(E == 128 && nd <= 15 && H % 128 == 0 && H <= 2048,
                "tensor-core rollout needs E=128, noise_dim<=15, H multiple of 128 and <=2048 (use precision fp32 otherwise)");
  const int nch = H / 128;
  TcPlan* P = new TcPlan();
  memset(P, 0, sizeof(*P));
  P->np = np; P->nd = nd; P->H = H; P->n_chunks = nch; P->W1 = Wt[0]; P->b1 = bt[0];
  P->gext_stride = 128 * 32 * (np == 3 ? 2 : 1);
  const size_t cap = (size_t)(np == 3 ? 2 : 1) * ((size_t)(128 * 144 + 64 * 144 + 128 * 80 + 256 * 144 + 16 * 272) * 2 +
                                                 (size_t)nch * (128 * 144 + 128 * 144) * 2) + 128 * kTcMaxItems;
  cudaError_t e = cudaMalloc(&P->stream, cap);
  if (e == cudaSuccess) e = cudaMalloc(&P->items, sizeof(TcItem) * kTcMaxItems);
  if (e != cudaSuccess) {
    set_error("tensor-core plan: cudaMalloc failed: %s", cudaGetErrorString(e));
    tc_plan_destroy(P);
    return NDP_ERR_CUDA;
  }
  cudaMemsetAsync(P->stream, 0, cap, s);
  Builder B;
  B.np = np; B.s = s; B.stream = P->stream;
  const int C1 = 2 * E + nd;
  ExtCols x;
  // generator fc1: state columns from TMEM (AS); ext = [noise cols | per-goal bias] built per call
  B.add(Wt[0], C1, 0, 128, 0, 128, 0, 4, kColR0, kColAS, 64, true, true, false, 0, nullptr);
  B.add(Wt[0], C1, 0, 128, 0, 128, 4, 4, kColR0, kColAS, 64, false, false, true, 1, nullptr);
  // fc2 (N=64), fc3 (K=64)
  x = bias_only(bt[1]);
  B.add(Wt[1], 128, 0, 64, 0, 64, 0, 8, kColR1, kColR0, 64, true, true, true, 3, &x);
  x = bias_only(bt[2]);
  B.add(Wt[2], 64, 0, 128, 0, 128, 0, 4, kColR0, kColR1, 32, true, true, true, 3, &x);
  // fc4 as two N=128 halves (R1, DR)
  for (int hh = 0; hh < 2; ++hh) {
    x = bias_only(bt[3] + hh * 128);
    B.add(Wt[3], 128, hh * 128, 128, 0, 128, 0, 4, hh ? kColDR : kColR1, kColR0, 64, true, hh == 0, false, 0, nullptr);
    B.add(Wt[3], 128, hh * 128, 128, 0, 128, 4, 4, hh ? kColDR : kColR1, kColR0, 64, false, false, true, 3, &x);
  }
  // fc5 padded to N=16 (4 real rows), K=256 spanning R1|DR
  x = bias_only(bt[4]);
  B.add(Wt[4], 256, 0, 4, 0, 16, 0, 16, kColR0, kColR1, 64, true, true, true, 3, &x);
  // forward model: F1[c] = fc1 rows of hidden chunk c (ext = [action cols | bias]); F2[c] = fc2 columns of chunk c
  auto F1 = [&](int c) {
    ExtCols a;
    memset(&a, 0, sizeof(a));
    for (int k = 0; k < 4; ++k) { a.p[k] = Wt[5] + (size_t)c * 128 * (E + 4) + E + k; a.stride[k] = E + 4; }
    a.p[4] = bt[5] + c * 128; a.stride[4] = 1;
    const int dcol = (c & 1) ? kColR1 : kColR0;
    B.add(Wt[5], E + 4, c * 128, 128, 0, 128, 0, 4, dcol, kColAS, 64, true, c == 0, false, 0, nullptr);
    B.add(Wt[5], E + 4, c * 128, 128, 0, 128, 4, 4, dcol, kColAS, 64, false, false, true, 2, &a);
  };
  auto F2 = [&](int c) {
    const int acol = (c & 1) ? kColR1 : kColR0;
    const bool last = (c == nch - 1);
    ExtCols b = bias_only(bt[6]);
    B.add(Wt[6], H, 0, 128, c * 128, 128, 0, 4, kColDR, acol, 64, c == 0, true, false, 0, nullptr);
    B.add(Wt[6], H, 0, 128, c * 128, 128, 4, 4, kColDR, acol, 64, false, false, last, last ? 3 : 0, last ? &b : nullptr);
  };
  F1(0);
  for (int c = 1; c < nch; ++c) { F1(c); F2(c - 1); }
  F2(nch - 1);
  if (B.rc) { tc_plan_destroy(P); return B.rc; }
  P->n_items = B.n;
  P->stream_bytes = B.off;
  e = cudaMemcpyAsync(P->items, B.items, sizeof(TcItem) * B.n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);   // B.items is on this stack frame
  if (e != cudaSuccess) {
    set_error("tensor-core plan: upload failed: %s", cudaGetErrorString(e));
    tc_plan_destroy(P);
    return NDP_ERR_CUDA;
  }
  static const bool one_tile = getenv("NDP_TC_ONE_TILE") != nullptr;     // diagnostics: keep the one-tile kernel
  if (np == 1 && !one_tile && H % 128 == 0)
    if (int rc2 = tc2_build(P, E, nd, H, Wt, bt, s)) { tc_plan_destroy(P); return rc2; }
  *out = P;
  return NDP_OK;
}

// ==========================================================================================================
// Two-tile bulk kernel (NP = 1, the single-pass estimate of "f16" / "f16+refine"): TWO 128-candidate tiles of one
// goal per CTA, 256 TMEM columns each, their MMA bursts alternating in the tensor queue.
//
// Why: with one tile per SM the tensor pipe idles ~40 % of the time behind the strictly serial layer chain (every
// layer's MMAs wait for the previous layer's TMEM -> ReLU -> fp16 -> TMEM epilogue; the round trip costs ~1100 clk in
// this kernel, tools/tc2_prof.py).  Two independent tiles hide each other's round
// trips as long as one tile's MMA burst is about as long as the other's round trip, so every event is made as large
// as TMEM allows: N = 128 everywhere but fc2 / fc5, and 128-wide hidden chunks in the forward model.  That needs the
// state operand OUT of TMEM: the state lives in shared memory as an fp16 pair hi + lo (22 significand bits; the
// per-step rounding of 2^-22 is far below the fp16 rounding of the activations), and `hi` IS the MMA A operand
// (canonical K-major core-matrix layout, SS-mode MMAs for generator fc1 and forward fc1) — same 64 KB per tile as an
// fp32 state, no TMEM columns.  Both tiles consume the SAME weight ring slot (released after the second tile's MMAs), so the L2 -> shared-memory weight traffic per candidate halves.
//
// TMEM per tile (base 256 * tile):  P [0,128)  Q [128,256)
//   G1  D=P <- state(smem)             epilogue P -> Q[0,64)
//   G2  D=P[0,64) <- Q[0,64)           epilogue     -> Q[64,96)
//   G3  D=P <- Q[64,96)                epilogue P -> Q[0,64)
//   G4a D=P <- Q[0,64)                 epilogue P -> Q[64,128)      (fc4 outputs 0..127)
//   G4b D=P <- Q[0,64)                 epilogue P -> Q[0,64)        (fc4 outputs 128..255)
//   G5  D=P[0,16) <- Q[64,128) | Q[0,64)
//   F1[c] (hidden chunk of 128)  D=Q <- state(smem), epilogue in place;  F2[c]  DR=P += Q
// Products and accumulation order per element equal rollout_tc_kernel<1>; only the state is carried as fp16 hi + 8-bit lo (14 bits) instead of fp32.
// ==========================================================================================================
static constexpr int kT2Threads = 576;                   // 16 epilogue warps, MMA issuer (16), TMA producer (17)
static constexpr uint32_t kT2P = 0, kT2Q = 128;
// per tile: hi image 32 KB (fp16, A operand layout) | lo image 16 KB.  The lo image is private to the epilogue threads
// (no MMA reads it): it only carries the state's rounding residual s - hi from one step's END to the next, so it is kept
// in 8 bits per element — the high byte of fp16(s - hi), i.e. e5m2 with round-to-nearest: the state is carried at
// 11 + 3 bits instead of 11 + 11, an error of <= 2^-14 |s| per step against the 2^-12 |s| operand rounding of every product
// of this single-pass estimate.  The 32 KB this frees hold a fifth ring slot (ring waits 4.6 k -> 3.3 k cycles per step).
// lo element (row m, k): byte (k/8)*1024 + m*8 + k%8.
static constexpr uint32_t kT2OffState = 0;
static constexpr uint32_t kT2TileState = 49152, kT2LoBase = 32768, kT2StateBytes = 2 * kT2TileState;
static constexpr uint32_t kT2NSlot = 5;                   // one GEMM (2 items) in use + one in flight + one spare
static constexpr uint32_t kT2OffExt = kT2StateBytes;     // noise[2] 2 x 4 KB | action[2] 2 x 2 KB | ones 2 KB | zeros 2 KB
static constexpr uint32_t kT2ExtBytes = 16384;           // (action / ones use k < 8 only: their second k-chunk is the shared zero block)
static constexpr uint32_t kT2OffRing = kT2OffExt + kT2ExtBytes;
static constexpr uint32_t kT2Slot = 20480;
// barrier indices (8 bytes each): full[NSlot] | empty[NSlot] | d_full[tile][2] | a_ready[tile][2]
static constexpr uint32_t kT2BarEmpty = kT2NSlot, kT2BarD = 2 * kT2NSlot, kT2BarA = 2 * kT2NSlot + 4;
static constexpr uint32_t kT2OffItems = kT2OffRing + kT2NSlot * kT2Slot;
static constexpr uint32_t kT2ItemsBytes = kTcMaxItems * 12;      // {src_off, bytes, ext_bytes} per item
static constexpr uint32_t kT2OffGoal = kT2OffItems + kT2ItemsBytes;
static constexpr uint32_t kT2OffRed = kT2OffGoal + 512;          // 2 x 128 floats
static constexpr uint32_t kT2OffBars = kT2OffRed + 1024;
static constexpr uint32_t kT2SmemBytes = kT2OffBars + 256;
static_assert(kT2SmemBytes <= 232448, "two-tile kernel: shared memory map exceeds 227 KB");

struct T2Tile {              // smem operand descriptors of one tile: (address >> 4) | LBO bits
  uint32_t ext_noise, ext_act, ext_ones;
  uint32_t state16;          // state hi image: A operand of the SS-mode MMAs
};

// MMAs of one weight group: NMAIN main slabs + optional ext slab (A from smem).
// AMODE 0: A from TMEM at a + (s/SPC)*CPH + (s%SPC)*8;  AMODE 1: A = state hi image in smem, slabs a .. a+NMAIN-1
template <int N, int NMAIN, int EXT, int AMODE, int SPC, int CPH>
__device__ __forceinline__ void t2_group(const T2Tile& st, uint32_t b16, uint32_t d, uint32_t a, uint32_t first) {
  constexpr uint32_t kDescHi = 0x4000u | (128u >> 4);
  constexpr uint32_t kIdesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  constexpr uint32_t step = 2 * N;                              // one slab = N*32 bytes, in 16-byte units
  const uint32_t blo = b16 | ((uint32_t)N << 16);               // LBO = N*16 bytes
  uint32_t acc = first ? 0u : 1u;
#pragma unroll
  for (int s = 0; s < NMAIN; ++s) {
    const uint64_t bd = pack64(blo + s * step, kDescHi);
    if (AMODE == 0) mma_ts(d, a + (s / SPC) * CPH + (s % SPC) * 8, bd, kIdesc, acc);
    else mma_ss(d, pack64(st.state16 + (a + s) * 256, kDescHi), bd, kIdesc, acc);     // 4 KB per K=16 slab
    acc = 1;
  }
  if (EXT) {
    const uint32_t ea = EXT == 1 ? st.ext_noise : (EXT == 2 ? st.ext_act : st.ext_ones);
    mma_ss(d, pack64(ea, kDescHi), pack64(blo + NMAIN * step, kDescHi), kIdesc, acc);
  }
}

#ifdef NDP_TC_PROF
__device__ unsigned long long g_t2_prof[8];   // issuer: [0] wait_a [1] wait_full [2] total; epilogue warp 0: [4] wait_d [5] total
#endif

// relu fused into the fp16 pack: ONE instruction per pair (cvt.rn.relu.f16x2.f32) instead of 2 FMNMX + cvt.
// Same bits as pack_hi(max(x0, 0), max(x1, 0)) — rounding to nearest is monotonic and keeps the sign — except that a NaN
// stays a NaN (torch's relu keeps it too).
__device__ __forceinline__ uint32_t relu_pack(uint32_t x0, uint32_t x1) {
  uint32_t d;                                                   // one F2FP with the relu modifier (x1 -> upper half)
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(__uint_as_float(x1)), "f"(__uint_as_float(x0)));
  return d;
}
// 64 accumulator columns (src, src+32) -> relu -> 32 packed fp16 columns (dst, dst+16); both loads in flight at once
__device__ __forceinline__ void t2_relu64(uint32_t src, uint32_t dst) {
  uint32_t v[32], w[32];
  tmem_ld32(src, v);
  tmem_ld32(src + 32, w);
  tmem_ld_wait();
  uint32_t hi[16];
#pragma unroll
  for (int j = 0; j < 16; ++j)
    hi[j] = relu_pack(v[2 * j], v[2 * j + 1]);
  tmem_st16(dst, hi);
#pragma unroll
  for (int j = 0; j < 16; ++j)
    hi[j] = relu_pack(w[2 * j], w[2 * j + 1]);
  tmem_st16(dst + 16, hi);
}
// 32 accumulator columns -> relu -> 16 packed fp16 columns
__device__ __forceinline__ void t2_relu32(uint32_t src, uint32_t dst) {
  uint32_t v[32];
  tmem_ld32(src, v);
  tmem_ld_wait();
  uint32_t hi[16];
#pragma unroll
  for (int j = 0; j < 16; ++j)
    hi[j] = relu_pack(v[2 * j], v[2 * j + 1]);
  tmem_st16(dst, hi);
}

// CL = 1 (default): every CTA streams the weights on its own.  CL = 2 (NDP_TC2_CLUSTER=1, measured slower, kept as a
// reproducible negative result): the two CTAs of a thread-block cluster share the stream —
// each fetches HALF of every ring item and multicasts it into both CTAs' shared memory (data and mbarrier signal land
// at the same CTA-relative offsets), so the L2 -> SM traffic of the weight stream (24.6 B/clk/SM = 6 TB/s chip-wide at
// cfg2, the reason the issuer waited 14 % of its time for ring slots) is halved.  A slot is refilled once BOTH CTAs have
// released it (the issuers' tcgen05.commit arrives on the empty barrier of both CTAs, barrier count CL).
// hand-over of a split pair's state images (48 KB per tile) through global memory.  __noinline__: kept out of the
// kernel body so that the step loops of the SPLIT instantiation are scheduled / register-allocated like the lean one
__device__ __forceinline__ void t2_tile_bar(int tile) {
  if (tile == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
  else asm volatile("bar.sync 2, 256;" ::: "memory");
}
__device__ __noinline__ void t2_state_restore(uint8_t* tile_smem, const uint8_t* src_g, const unsigned* flag, int tt, int tile) {
  if (tt == 0) {                                            // bounded wait for the first piece of this tile
    const long long c0 = clock64();
    while (ld_acquire_gpu(flag) == 0u && clock64() - c0 < 4000000000ll) __nanosleep(128);
  }
  t2_tile_bar(tile);
  const uint4* src = reinterpret_cast<const uint4*>(src_g);
  uint4* dst = reinterpret_cast<uint4*>(tile_smem);
#pragma unroll 4
  for (int j = 0; j < (int)(kT2TileState / 4096); ++j) dst[tt + 256 * j] = __ldcg(src + tt + 256 * j);
}
__device__ __noinline__ void t2_state_dump(const uint8_t* tile_smem, uint8_t* dst_g, unsigned* flag, int tt, int tile) {
  t2_tile_bar(tile);                                        // every thread's put_state8 of the last step
  const uint4* src = reinterpret_cast<const uint4*>(tile_smem);
  uint4* dst = reinterpret_cast<uint4*>(dst_g);
#pragma unroll 4
  for (int j = 0; j < (int)(kT2TileState / 4096); ++j) dst[tt + 256 * j] = src[tt + 256 * j];
  __threadfence();
  t2_tile_bar(tile);
  if (tt == 0) st_release_gpu(flag, 1u);
}

// SPLIT = false: the launch has no time-split pairs (every pipelined PlanRing launch) — none of the job mapping /
// hand-over code is compiled in (with it, the un-split kernel ran 2.5 % slower in the pipelined loop: same-box A/B of
// the round-2b and round-2e builds, 1.757 vs 1.712e9 sustained).
template <int CL, bool SPLIT>
__global__ void __launch_bounds__(kT2Threads, 1) rollout_tc2_kernel(const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t crank = CL > 1 ? cluster_ctarank() : 0u;
  constexpr uint16_t kClusterMask = (uint16_t)((1u << CL) - 1u);
  const bool ghost = CL > 1 && (int)blockIdx.x >= p.n_ctas;
#ifdef NDP_TC_PROF
  // whole-CTA life time in SM clocks [7] and in nanoseconds [6] (their ratio is the effective SM clock), CTA count [3]
  const long long t2p_cta0 = clock64();
  unsigned long long t2p_ns0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2p_ns0));
#endif
  // ---- job of this CTA: a tile pair and a step range [t0, t1).  Without a split (p.sp_n == 0) CTA b rolls pair b over
  // all T steps.  With one, the pairs of the last two waves are laid out so that the SMs finish together instead of
  // leaving the last wave partly empty (t2_split() below): CTAs are dispatched in index order, one per SM —
  //   [base full pairs] [A: n full pairs] [n first pieces, steps [0,a)] [H full pairs] [n continuations, steps [a,T)]
  // the H "helper" SMs run the first pieces of the n split pairs BEFORE their own full pair; the n other SMs append
  // a continuation to theirs.  A first piece leaves the pair's state images (the fp16 hi/lo split the next step
  // would read anyway: bit-identical to the unsplit roll) in global scratch and raises a flag per tile.
  int job = ghost ? p.n_ctas - 1 : (int)blockIdx.x, t0 = 0, t1 = p.T, sp_id = -1;
  if (SPLIT && p.sp_n > 0 && job >= p.sp_base) {
    const int b = job - p.sp_base, n = p.sp_n;
    if (b < n) job = p.sp_base + b;
    else if (b < 2 * n) { sp_id = b - n; job = p.sp_base + n + sp_id; t1 = p.sp_a; }
    else if (b < 2 * n + p.sp_h) job = p.sp_base + 2 * n + (b - 2 * n);
    else { sp_id = b - 2 * n - p.sp_h; job = p.sp_base + n + sp_id; t0 = p.sp_a; }
  }
  const int g = job / p.tiles_per_goal;                         // tiles_per_goal counts tile PAIRS here
  const int pair = job % p.tiles_per_goal;

  uint8_t* ext_base = smem + kT2OffExt;                         // noise0, noise1 (4 KB each), act0, act1, ones, zeros (2 KB each)
  uint8_t* ext_actb = ext_base + 2 * kExtSlab;
  uint8_t* ext_ones = ext_actb + 4096;
  uint8_t* ext_zero = ext_ones + 2048;
  uint32_t* items = reinterpret_cast<uint32_t*>(smem + kT2OffItems);   // [n_items][3]: src_off, bytes, ext_bytes
  float* s_goal = reinterpret_cast<float*>(smem + kT2OffGoal);
  float* s_red = reinterpret_cast<float*>(smem + kT2OffRed);
  const uint32_t bars = smem_u32(smem + kT2OffBars);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kT2OffBars + 192);

  for (int i = tid; i < p.n_items * 3; i += kT2Threads)
    items[i] = reinterpret_cast<const uint32_t*>(p.items)[(i / 3) * 8 + i % 3];      // TcItem: src_off, bytes, ext_bytes first
  for (int i = tid; i < (int)kT2ExtBytes / 16; i += kT2Threads)
    reinterpret_cast<uint4*>(ext_base)[i] = make_uint4(0, 0, 0, 0);
  if (tid < 128) s_goal[tid] = __ldg(p.goal + (size_t)g * 128 + tid);
  if (tid == 0) {
    // barrier map (8 bytes each): full[0..3] empty[4..7] d_full[tile][2] @8..11  a_ready[tile][2] @12..15
    for (uint32_t s = 0; s < kT2NSlot; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (kT2BarEmpty + s), CL); }
    for (uint32_t e = 0; e < 4; ++e) { mbar_init(bars + 8 * (kT2BarD + e), 1); mbar_init(bars + 8 * (kT2BarA + e), 8); }
    fence_mbar_init();
  }
  if (warp == 16) tmem_alloc512(smem_u32((const void*)tmem_slot));
  __syncthreads();
  if (tid < 128) *reinterpret_cast<uint16_t*>(ext_ones + (tid >> 3) * 128 + (tid & 7) * 16) = 0x3C00;   // k=0 column = 1.0
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  if (CL > 1) cluster_sync_all();                              // the peer's barriers exist before anything is sent to them
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp == 17) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      const uint8_t* gsrc = p.gext + (size_t)g * p.gext_stride;
      for (int t = t0; t < t1; ++t) {
        for (int i = 0; i < p.n_items; ++i, ++it) {
          const uint32_t slot = it % kT2NSlot, par = (it / kT2NSlot) & 1;
          const uint32_t bytes = items[3 * i + 1], eb = items[3 * i + 2];
          mbar_wait(bars + 8 * (kT2BarEmpty + slot), par ^ 1);
          mbar_expect_tx(bars + 8 * slot, bytes + eb);
          const uint32_t dst = smem_u32(smem + kT2OffRing + slot * kT2Slot);
          if (CL == 1 || (p.dbg & 1)) {   // (CL == 1: decided at compile time)
            bulk_g2s(dst, p.stream + items[3 * i], bytes, bars + 8 * slot);
          } else {                                          // my 1/CL of the item, to every CTA of the cluster
            const uint32_t part = bytes / CL, o = crank * part;
            bulk_g2s_mc(dst + o, p.stream + items[3 * i] + o, part, bars + 8 * slot, kClusterMask);
          }
          if (eb) bulk_g2s(dst + bytes, gsrc, eb, bars + 8 * slot);   // per-goal block: this CTA only
        }
      }
    }
  } else if (warp == 16) {
    // ===================== MMA issuer (one warp, both tiles) =====================
    // Per GEMM (a "section" of 1-2 ring items): tile 0's MMAs, then tile 1's, back to back in the tensor queue, so
    // tile 0's epilogue round trip runs under tile 1's MMAs and vice versa.  (Two concurrently issuing warps would
    // interleave the tiles' MMAs in the queue: both bursts would complete together and both round trips would idle the
    // pipe at the same time — measured, tools/tc2_prof.py.)  A ring slot is released after tile 1's MMAs on it.
    T2Tile td[2];
    const uint32_t ext_lbo = (2048u >> 4) << 16;               // smem A operands: LBO = 2 KB, SBO = 128 B
    // descriptor start-address field: 14 bits of (address >> 4).  Inside a thread-block cluster the shared::cta
    // addresses of a CTA carry its window's high bits; unmasked they would spill into the LBO field (rank 1 of a
    // cluster computed garbage until these were masked)
    auto a16 = [](uint32_t saddr) { return (saddr >> 4) & 0x3FFFu; };
#pragma unroll
    for (int tl = 0; tl < 2; ++tl) {
      td[tl].ext_noise = a16(smem_u32(ext_base + tl * kExtSlab)) | ext_lbo;
      // action / ones slabs hold k-chunk 0 only; LBO points their k-chunk 1 at the shared zero block
      td[tl].ext_act = a16(smem_u32(ext_actb + tl * 2048)) | (((smem_u32(ext_zero) - smem_u32(ext_actb + tl * 2048)) >> 4) << 16);
      td[tl].ext_ones = a16(smem_u32(ext_ones)) | (((smem_u32(ext_zero) - smem_u32(ext_ones)) >> 4) << 16);
      td[tl].state16 = a16(smem_u32(smem + kT2OffState + tl * kT2TileState)) | ext_lbo;
    }
    const uint32_t ring16 = a16(smem_u32(smem + kT2OffRing));
    uint32_t slot = 0, par = 0, ds[2] = {0, 0}, aw[2] = {0, 0};
    const int nch = p.n_hid_chunks;                             // hidden chunks of 128
#ifdef NDP_TC_PROF
    long long t2p_a = 0, t2p_f = 0;
    const long long t2p_begin = clock64();
#endif
    auto wait_a = [&](int tl) {
#ifdef NDP_TC_PROF
      const long long c0 = clock64();
#endif
      mbar_wait(bars + 8 * (kT2BarA + 2 * tl + (aw[tl] & 1)), (aw[tl] >> 1) & 1);
      ++aw[tl];
#ifdef NDP_TC_PROF
      t2p_a += clock64() - c0;
#endif
    };
    auto wait_full = [&](uint32_t j) {                          // ring item `j` items ahead of the section start
#ifdef NDP_TC_PROF
      const long long c0 = clock64();
#endif
      const uint32_t sj = slot + j;
      mbar_wait(bars + 8 * (sj >= kT2NSlot ? sj - kT2NSlot : sj), sj >= kT2NSlot ? (par ^ 1) : par);
#ifdef NDP_TC_PROF
      t2p_f += clock64() - c0;
#endif
    };
    auto slot16 = [&](uint32_t j) { const uint32_t sj = slot + j; return ring16 + (sj >= kT2NSlot ? sj - kT2NSlot : sj) * (kT2Slot >> 4); };
    auto free_slot = [&](uint32_t j) {
      const uint32_t sj = slot + j, b = bars + 8 * (kT2BarEmpty + (sj >= kT2NSlot ? sj - kT2NSlot : sj));
      if (CL == 1) mma_commit(b); else mma_commit_mc(b, kClusterMask);
    };
    auto signal_d = [&](int tl) { mma_commit(bars + 8 * (kT2BarD + 2 * tl + (ds[tl] & 1))); };
    auto advance = [&](uint32_t n) { slot += n; if (slot >= kT2NSlot) { slot -= kT2NSlot; par ^= 1; } };
    // a GEMM of two ring items: slabs 0..3 (item 0), slabs 4..7 + ext (item 1)
    // AM: 1 = A from the state image (slabs a, a+4), 0 = A from TMEM (columns a, a + a_hi)
#define T2_GEMM2(N_, EXT_, AM_, DOFF, A0, A1, FIRST, WAIT, SIGNAL)                                           \
    for (int tl = 0; tl < 2; ++tl) {                                                                        \
      const uint32_t tb_ = tmem + 256 * tl;                                                                 \
      if (WAIT) wait_a(tl);                                                                                 \
      if (tl == 0) wait_full(0);                                                                            \
      fence_after_sync();                                                                                   \
      if (elect_one()) {                                                                                    \
        t2_group<N_, 4, 0, AM_, 16, 0>(td[tl], slot16(0), tb_ + (DOFF), (AM_) ? (A0) : tb_ + (A0), FIRST);   \
        if (tl == 1) free_slot(0);                                                                          \
      }                                                                                                     \
      __syncwarp();                                                                                         \
      if (tl == 0) wait_full(1);                                                                            \
      fence_after_sync();                                                                                   \
      if (elect_one()) {                                                                                    \
        t2_group<N_, 4, EXT_, AM_, 16, 0>(td[tl], slot16(1), tb_ + (DOFF), (AM_) ? (A1) : tb_ + (A1), 0);    \
        if (tl == 1) free_slot(1);                                                                          \
        if (SIGNAL) signal_d(tl);                                                                           \
      }                                                                                                     \
      __syncwarp();                                                                                         \
      if (SIGNAL) ++ds[tl];                                                                                 \
    }                                                                                                       \
    advance(2);
    for (int t = t0; t < t1; ++t) {
      // ---- generator fc1: state slabs 0..3, then 4..7 + [noise | per-goal bias]  -> P
      T2_GEMM2(128, 1, 1, kT2P, 0, 4, 1, true, true)
      // ---- fc2 (N = 64, one item): A1 = Q[0,64) -> P[0,64)
      for (int tl = 0; tl < 2; ++tl) {
        const uint32_t tb_ = tmem + 256 * tl;
        wait_a(tl);
        if (tl == 0) wait_full(0);
        fence_after_sync();
        if (elect_one()) {
          t2_group<64, 8, 3, 0, 16, 0>(td[tl], slot16(0), tb_ + kT2P, tb_ + kT2Q, 1);
          if (tl == 1) free_slot(0);
          signal_d(tl);
        }
        __syncwarp();
        ++ds[tl];
      }
      advance(1);
      // ---- fc3 (K = 64, one item): A2 = Q[64,96) -> P
      for (int tl = 0; tl < 2; ++tl) {
        const uint32_t tb_ = tmem + 256 * tl;
        wait_a(tl);
        if (tl == 0) wait_full(0);
        fence_after_sync();
        if (elect_one()) {
          t2_group<128, 4, 3, 0, 16, 0>(td[tl], slot16(0), tb_ + kT2P, tb_ + kT2Q + 64, 1);
          if (tl == 1) free_slot(0);
          signal_d(tl);
        }
        __syncwarp();
        ++ds[tl];
      }
      advance(1);
      // ---- fc4 halves (outputs 0..127, 128..255): A3 = Q[0,64) -> P
      T2_GEMM2(128, 3, 0, kT2P, kT2Q, kT2Q + 32, 1, true, true)
      T2_GEMM2(128, 3, 0, kT2P, kT2Q, kT2Q + 32, 1, true, true)
      // ---- fc5 (N = 16, K = 256, one item): A4a = Q[64,128) (k 0..127), A4b = Q[0,64) (k 128..255) -> P[0,16)
      for (int tl = 0; tl < 2; ++tl) {
        const uint32_t tb_ = tmem + 256 * tl;
        wait_a(tl);
        if (tl == 0) wait_full(0);
        fence_after_sync();
        if (elect_one()) {
          t2_group<16, 8, 0, 0, 16, 0>(td[tl], slot16(0), tb_ + kT2P, tb_ + kT2Q + 64, 1);
          t2_group<16, 8, 3, 0, 16, 0>(td[tl], slot16(0) + (4096 >> 4), tb_ + kT2P, tb_ + kT2Q, 0);
          if (tl == 1) free_slot(0);
          signal_d(tl);
        }
        __syncwarp();
        ++ds[tl];
      }
      advance(1);
      // ---- forward model: F1[0] -> Q; then per hidden chunk F2[c] (DR = P += Q), F1[c+1] -> Q
      T2_GEMM2(128, 2, 1, kT2Q, 0, 4, 1, true, true)
      // per hidden chunk a BURST of four ring items per tile: F2[c] (DR = P += Q) then F1[c+1] (-> Q); tile 0's whole
      // burst, then tile 1's: a tile's epilogue round trip then runs under the other tile's 1088 tensor cycles.
      // (All four ring slots hold the burst; slot j is released by tile 1 one burst before tile 0 needs item j + 4.)
#pragma unroll 1
      for (int c = 0; c < nch; ++c) {
        const uint32_t first = (c == 0) ? 1u : 0u;
        if (c == nch - 1) {
          T2_GEMM2(128, 3, 0, kT2P, kT2Q, kT2Q + 64, first, true, true)
        } else {
          for (int tl = 0; tl < 2; ++tl) {
            const uint32_t tb_ = tmem + 256 * tl;
            wait_a(tl);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (tl == 0) wait_full(j);
              fence_after_sync();
              if (elect_one()) {
                if (j == 0) t2_group<128, 4, 0, 0, 16, 0>(td[tl], slot16(0), tb_ + kT2P, tb_ + kT2Q, first);
                if (j == 1) t2_group<128, 4, 0, 0, 16, 0>(td[tl], slot16(1), tb_ + kT2P, tb_ + kT2Q + 64, 0);
                if (j == 2) t2_group<128, 4, 0, 1, 16, 0>(td[tl], slot16(2), tb_ + kT2Q, 0, 1);
                if (j == 3) t2_group<128, 4, 2, 1, 16, 0>(td[tl], slot16(3), tb_ + kT2Q, 4, 0);
                if (tl == 1) free_slot(j);
                if (j == 3) signal_d(tl);
              }
              __syncwarp();
            }
            ++ds[tl];
          }
          advance(4);
        }
      }
    }
#undef T2_GEMM2
#ifdef NDP_TC_PROF
    if (lane == 0) {
      atomicAdd(&g_t2_prof[0], (unsigned long long)t2p_a);
      atomicAdd(&g_t2_prof[1], (unsigned long long)t2p_f);
      atomicAdd(&g_t2_prof[2], (unsigned long long)(clock64() - t2p_begin));
    }
#endif
  } else {
    // ===================== epilogue warps: tile = warp / 8 =====================
    const int tile = warp >> 3, q = warp & 3, h = (warp >> 2) & 1;
    const int m = q * 32 + lane;                                // candidate row = TMEM lane
    const int cand = (pair * 2 + tile) * 128 + m;
    const bool valid = cand < p.K && !ghost;
    // state images of this tile: element (row m, k) at (k/8)*2048 + (m/8)*128 + (m%8)*16 + (k%8)*2
    const uint32_t rowoff = (uint32_t)(m >> 3) * 128 + (uint32_t)(m & 7) * 16;
    uint8_t* SH = smem + kT2OffState + tile * kT2TileState + rowoff;
    uint8_t* SL = smem + kT2OffState + tile * kT2TileState + kT2LoBase + m * 8;      // 8-bit lo image: 8 bytes per (k-chunk, row)
    uint8_t* ext_noise = ext_base + tile * kExtSlab;
    uint8_t* ext_act = ext_actb + tile * 2048;
    const uint32_t lane_base = tmem + 256 * tile + ((uint32_t)(q * 32) << 16);
    const uint32_t P = lane_base + kT2P, Q = lane_base + kT2Q;
    const uint32_t bar_d = bars + 8 * (kT2BarD + 2 * tile), bar_a = bars + 8 * (kT2BarA + 2 * tile);
    uint32_t dw = 0, ae = 0;
#ifdef NDP_TC_PROF
    long long t2p_d = 0;
    const long long t2p_e0 = clock64();
#endif
    auto wait_d = [&]() {
#ifdef NDP_TC_PROF
      const long long c0 = clock64();
#endif
      mbar_wait(bar_d + 8 * (dw & 1), (dw >> 1) & 1);
      ++dw;
      fence_after_sync();
#ifdef NDP_TC_PROF
      t2p_d += clock64() - c0;
#endif
    };
    auto signal_a = [&](bool wrote_smem = false) {
      tmem_st_wait();
      if (wrote_smem) fence_proxy_async();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a + 8 * (ae & 1));
      ++ae;
    };
    auto noise_row = [&](int t) {                               // ext row [noise.., 1] of step t (h == 0 threads)
      float x[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) x[k] = 0.f;
      if (t < p.T) {
        const float* src = p.noise + (((size_t)t * p.G + g) * p.K + cand) * p.nd;
#pragma unroll
        for (int k = 0; k < 15; ++k)
          if (k < p.nd && valid) x[k] = __ldg(src + k);
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (k == p.nd) x[k] = 1.f;
      }
      const uint32_t off = (uint32_t)(m >> 3) * 128 + (uint32_t)(m & 7) * 16;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c == 1 && p.nd < 8) break;
        uint4 hv;
        hv.x = pack_hi(x[8 * c + 0], x[8 * c + 1]); hv.y = pack_hi(x[8 * c + 2], x[8 * c + 3]);
        hv.z = pack_hi(x[8 * c + 4], x[8 * c + 5]); hv.w = pack_hi(x[8 * c + 6], x[8 * c + 7]);
        *reinterpret_cast<uint4*>(ext_noise + c * 2048 + off) = hv;
      }
    };
    // noise_dim <= 4 (the reference's is 2): the NEXT step's noise is fetched into registers at the top of the step, so
    // that the END phase (on the step's critical path: the issuer waits for it before the next G1) does not expose
    // the global-load latency
    const bool npf = p.nd <= 4;
    float npre[4] = {0.f, 0.f, 0.f, 0.f};
    auto noise_prefetch = [&](int t) {
      if (t < p.T && valid) {
        const float* src = p.noise + (((size_t)t * p.G + g) * p.K + cand) * p.nd;
#pragma unroll
        for (int k = 0; k < 4; ++k) npre[k] = k < p.nd ? __ldg(src + k) : 0.f;
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) npre[k] = 0.f;
      }
    };
    auto noise_row_pre = [&](int t) {                           // == noise_row(t) from the prefetched values
      float x[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = k < 4 ? npre[k] : 0.f;
      if (t < p.T) {
#pragma unroll
        for (int k = 0; k < 5; ++k)
          if (k == p.nd) x[k] = 1.f;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = 0.f;
      }
      uint4 hv;
      hv.x = pack_hi(x[0], x[1]); hv.y = pack_hi(x[2], x[3]); hv.z = pack_hi(x[4], x[5]); hv.w = pack_hi(x[6], x[7]);
      *reinterpret_cast<uint4*>(ext_noise + (uint32_t)(m >> 3) * 128 + (uint32_t)(m & 7) * 16) = hv;
    };
    // 8 consecutive state values (k = 8*kc .. 8*kc+7) -> hi / lo images
    auto put_state8 = [&](int kc, const float (&s)[8]) {
      uint4 hv, lv;
      hv.x = pack_hi(s[0], s[1]); lv.x = pack_lo(s[0], s[1], hv.x);
      hv.y = pack_hi(s[2], s[3]); lv.y = pack_lo(s[2], s[3], hv.y);
      hv.z = pack_hi(s[4], s[5]); lv.z = pack_lo(s[4], s[5], hv.z);
      hv.w = pack_hi(s[6], s[7]); lv.w = pack_lo(s[6], s[7], hv.w);
      *reinterpret_cast<uint4*>(SH + kc * 2048) = hv;
      // fp16 residuals -> their high bytes, rounded to nearest (+0x80 per half; residuals are far from the fp16 maximum)
      uint2 l8;
      l8.x = __byte_perm(lv.x + 0x00800080u, lv.y + 0x00800080u, 0x7531);
      l8.y = __byte_perm(lv.z + 0x00800080u, lv.w + 0x00800080u, 0x7531);
      *reinterpret_cast<uint2*>(SL + kc * 1024) = l8;
    };
    // ---- init: state <- e0 (this thread: k in [64h, 64h+64)) or the images a first piece left; noise ext of step t0
    const int tt = (warp & 7) * 32 + lane;                      // thread of the tile (256)
    if (!SPLIT || t0 == 0) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(p.e0 + (size_t)g * 128) + h * 16 + 2 * c);
        const float4 b = __ldg(reinterpret_cast<const float4*>(p.e0 + (size_t)g * 128) + h * 16 + 2 * c + 1);
        const float s8[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        put_state8(8 * h + c, s8);
      }
    } else {
      // continuation: wait (bounded) for the first piece of this tile, then copy its 64 KB of state images back
      t2_state_restore(smem + kT2OffState + tile * kT2TileState, p.sp_state + ((size_t)sp_id * 2 + tile) * 65536,
                       p.sp_flags + 2 * sp_id + tile, tt, tile);
    }
    if (h == 0) noise_row(t0);
    signal_a(true);
    for (int t = t0; t < t1; ++t) {
      if (h == 0 && npf) noise_prefetch(t + 1);
      // ---- G1 (P) -> Q[0,64) ; G2 (P[0,64)) -> Q[64,96) ; G3 (P) -> Q[0,64) ; G4a (P) -> Q[64,128) ; G4b (P) -> Q[0,64)
      wait_d(); t2_relu64(P + 64 * h, Q + 32 * h); signal_a();
      wait_d(); t2_relu32(P + 32 * h, Q + 64 + 16 * h); signal_a();
      wait_d(); t2_relu64(P + 64 * h, Q + 32 * h); signal_a();
      wait_d(); t2_relu64(P + 64 * h, Q + 64 + 32 * h); signal_a();
      wait_d(); t2_relu64(P + 64 * h, Q + 32 * h); signal_a();
      // ---- G5 (4 real columns in P[0,4)): action -> ext slab; step-0 actions are the plan's a0
      wait_d();
      if (h == 0) {
        uint32_t av[4];
        tmem_ld4(P, av);
        tmem_ld_wait();
        const float a0 = __uint_as_float(av[0]), a1 = __uint_as_float(av[1]), a2 = __uint_as_float(av[2]), a3 = __uint_as_float(av[3]);
        uint4 hv;
        hv.x = pack_hi(a0, a1); hv.y = pack_hi(a2, a3); hv.z = pack_hi(1.f, 0.f); hv.w = 0u;
        *reinterpret_cast<uint4*>(ext_act + (uint32_t)(m >> 3) * 128 + (uint32_t)(m & 7) * 16) = hv;
        if (t == 0 && valid)   // mpc_eval.py:150-151
          *reinterpret_cast<float4*>(p.a0 + ((size_t)g * p.K + cand) * 4) = make_float4(a0, a1, a2, a3);
      }
      signal_a(true);
      // ---- forward model hidden chunks of 128: F1[c] accumulator in Q -> relu -> A of F2[c], in place
#pragma unroll 1
      for (int c = 0; c < p.n_hid_chunks; ++c) { wait_d(); t2_relu64(Q + 64 * h, Q + 64 * h); signal_a(); }
      // ---- END: state += residual (forward_encoder.py:113-114), re-split into hi / lo, next noise
      wait_d();
      const bool more = t + 1 < p.T;
      float cs = 0.f;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        uint32_t r[32];
        tmem_ld32(P + 64 * h + 32 * u, r);
        uint4 hv[4];
        uint2 l8[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          hv[c] = *reinterpret_cast<const uint4*>(SH + (8 * h + 4 * u + c) * 2048);
          l8[c] = *reinterpret_cast<const uint2*>(SL + (8 * h + 4 * u + c) * 1024);
        }
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t hw[4] = {hv[c].x, hv[c].y, hv[c].z, hv[c].w};
          // bytes back into the high halves of fp16 pairs
          const uint32_t lw[4] = {__byte_perm(l8[c].x, 0u, 0x1404), __byte_perm(l8[c].x, 0u, 0x3424),
                                  __byte_perm(l8[c].y, 0u, 0x1404), __byte_perm(l8[c].y, 0u, 0x3424)};
          float s8[8];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&hw[e]));
            const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&lw[e]));
            s8[2 * e] = (fh.x + fl.x) + __uint_as_float(r[8 * c + 2 * e]);
            s8[2 * e + 1] = (fh.y + fl.y) + __uint_as_float(r[8 * c + 2 * e + 1]);
          }
          if (more) {
            put_state8(8 * h + 4 * u + c, s8);
          } else {
            // terminal cost: mean over E of squared error (nn.MSELoss, mpc_eval.py:161)
            const float* gl = s_goal + 64 * h + 32 * u + 8 * c;
#pragma unroll
            for (int e = 0; e < 8; ++e) { const float df = s8[e] - gl[e]; cs = fmaf(df, df, cs); }
            if (p.state_T && valid) {
              float4* o = reinterpret_cast<float4*>(p.state_T + ((size_t)g * p.K + cand) * 128 + 64 * h + 32 * u + 8 * c);
              o[0] = make_float4(s8[0], s8[1], s8[2], s8[3]);
              o[1] = make_float4(s8[4], s8[5], s8[6], s8[7]);
            }
          }
        }
      }
      if (more) {
        if (h == 0) { if (npf) noise_row_pre(t + 1); else noise_row(t + 1); }
        signal_a(true);
      } else {
        if (h == 1) s_red[tile * 128 + m] = cs;
        if (tile == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
        else asm volatile("bar.sync 2, 256;" ::: "memory");
        if (h == 0 && valid) p.cost[(size_t)g * p.K + cand] = (cs + s_red[tile * 128 + m]) * (1.0f / 128.0f);
      }
    }
    if (SPLIT && t1 < p.T && !ghost) {
      // first piece of a split pair: leave the state images for the continuation, then raise this tile's flag
      t2_state_dump(smem + kT2OffState + tile * kT2TileState, p.sp_state + ((size_t)sp_id * 2 + tile) * 65536,
                    p.sp_flags + 2 * sp_id + tile, tt, tile);
    }
#ifdef NDP_TC_PROF
    if (warp == 0 && lane == 0) {
      atomicAdd(&g_t2_prof[4], (unsigned long long)t2p_d);
      atomicAdd(&g_t2_prof[5], (unsigned long long)(clock64() - t2p_e0));
    }
#endif
  }
  fence_before_sync();
  __syncthreads();
  if (CL > 1) cluster_sync_all();        // no CTA leaves while its peer may still multicast into it or arrive on its barriers
  if (warp == 16) tmem_dealloc512(tmem);
#ifdef NDP_TC_PROF
  if (tid == 0) {
    unsigned long long ns1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns1));
    atomicAdd(&g_t2_prof[7], (unsigned long long)(clock64() - t2p_cta0));
    atomicAdd(&g_t2_prof[6], ns1 - t2p_ns0);
    atomicAdd(&g_t2_prof[3], 1ull);
  }
#endif
}

// host side: weight stream of the two-tile kernel (fp16 hi only), items in the kernel's consumption order
static int tc2_build(TcPlan* P, int E, int nd, int H, const float* const* Wt, const float* const* bt, cudaStream_t s) {
  const int nch = H / 128;
  P->n_chunks2 = nch;
  const size_t cap = (size_t)12 * 20480 + (size_t)nch * 4 * 20480 + 128 * kTcMaxItems;
  cudaError_t e = cudaMalloc(&P->stream2, cap);
  if (e == cudaSuccess) e = cudaMalloc(&P->items2, sizeof(TcItem) * kTcMaxItems);
  if (e != cudaSuccess) { set_error("two-tile plan: cudaMalloc failed: %s", cudaGetErrorString(e)); return NDP_ERR_CUDA; }
  cudaMemsetAsync(P->stream2, 0, cap, s);
  TcItem items[kTcMaxItems];
  int n = 0;
  size_t off = 0;
  const int C1 = 2 * E + nd;
  auto main_blk = [&](const float* W, int ld, int n0, int nvalid, int k0, int N, int nslab, size_t at) {
    const int total = N * nslab * 16;
    tc_pack_main_kernel<<<(total + 255) / 256, 256, 0, s>>>(reinterpret_cast<__half*>(P->stream2 + at), nullptr, W, ld, n0, k0, N, nslab, nvalid);
    count_launch();
    return (size_t)N * 32 * nslab;
  };
  auto ext_blk = [&](int N, int nvalid, const ExtCols& c, size_t at) {
    tc_pack_ext_kernel<<<(N * 16 + 255) / 256, 256, 0, s>>>(reinterpret_cast<__half*>(P->stream2 + at), nullptr, N, nvalid, c);
    count_launch();
    return (size_t)N * 32;
  };
  auto close_item = [&](size_t bytes, uint32_t ext_bytes) -> int {
    if (n >= kTcMaxItems) { set_error("two-tile plan: too many stream items (H too large)"); return NDP_ERR_ARG; }
    if (bytes + ext_bytes > kT2Slot) { set_error("two-tile plan: item of %zu bytes exceeds the ring slot", bytes + ext_bytes); return NDP_ERR_ARG; }
    TcItem& I = items[n++];
    memset(&I, 0, sizeof(I));
    I.src_off = (uint32_t)off; I.bytes = (uint32_t)bytes; I.ext_bytes = ext_bytes;
    off += (bytes + 127) / 128 * 128;
    return NDP_OK;
  };
  int rc;
  size_t b;
  ExtCols x;
  // G1: state columns 0..63 | 64..127 + per-goal ext block (from the workspace)
  b = main_blk(Wt[0], C1, 0, 128, 0, 128, 4, off);
  if ((rc = close_item(b, 0))) return rc;
  b = main_blk(Wt[0], C1, 0, 128, 64, 128, 4, off);
  if ((rc = close_item(b, 128 * 32))) return rc;
  // G2 (N=64), G3 (K=64)
  b = main_blk(Wt[1], 128, 0, 64, 0, 64, 8, off); x = bias_only(bt[1]); b += ext_blk(64, 64, x, off + b);
  if ((rc = close_item(b, 0))) return rc;
  b = main_blk(Wt[2], 64, 0, 128, 0, 128, 4, off); x = bias_only(bt[2]); b += ext_blk(128, 128, x, off + b);
  if ((rc = close_item(b, 0))) return rc;
  // G4 halves, two items each
  for (int hh = 0; hh < 2; ++hh) {
    b = main_blk(Wt[3], 128, 128 * hh, 128, 0, 128, 4, off);
    if ((rc = close_item(b, 0))) return rc;
    b = main_blk(Wt[3], 128, 128 * hh, 128, 64, 128, 4, off); x = bias_only(bt[3] + 128 * hh); b += ext_blk(128, 128, x, off + b);
    if ((rc = close_item(b, 0))) return rc;
  }
  // G5 padded to N=16 (4 real rows): k 0..127 | k 128..255 + bias
  b = main_blk(Wt[4], 256, 0, 4, 0, 16, 8, off);
  b += main_blk(Wt[4], 256, 0, 4, 128, 16, 8, off + b);
  x = bias_only(bt[4]); b += ext_blk(16, 4, x, off + b);
  if ((rc = close_item(b, 0))) return rc;
  // forward model: F1[c] = fc1 rows of hidden chunk c (ext = [action cols | bias]); F2[c] = fc2 columns of chunk c
  auto F1 = [&](int c) -> int {
    size_t bb = main_blk(Wt[5], E + 4, c * 128, 128, 0, 128, 4, off);
    if (int r = close_item(bb, 0)) return r;
    ExtCols a;
    memset(&a, 0, sizeof(a));
    for (int k = 0; k < 4; ++k) { a.p[k] = Wt[5] + (size_t)c * 128 * (E + 4) + E + k; a.stride[k] = E + 4; }
    a.p[4] = bt[5] + c * 128; a.stride[4] = 1;
    bb = main_blk(Wt[5], E + 4, c * 128, 128, 64, 128, 4, off);
    bb += ext_blk(128, 128, a, off + bb);
    return close_item(bb, 0);
  };
  auto F2 = [&](int c) -> int {
    size_t bb = main_blk(Wt[6], H, 0, 128, c * 128, 128, 4, off);
    if (int r = close_item(bb, 0)) return r;
    bb = main_blk(Wt[6], H, 0, 128, c * 128 + 64, 128, 4, off);
    if (c == nch - 1) { ExtCols bc = bias_only(bt[6]); bb += ext_blk(128, 128, bc, off + bb); }
    return close_item(bb, 0);
  };
  if ((rc = F1(0))) return rc;
  for (int c = 0; c < nch; ++c) {
    if ((rc = F2(c))) return rc;
    if (c + 1 < nch && (rc = F1(c + 1))) return rc;
  }
  if (cudaGetLastError() != cudaSuccess) { set_error("two-tile plan: pack kernel launch failed"); return NDP_ERR_CUDA; }
  P->n_items2 = n;
  e = cudaMemcpyAsync(P->items2, items, sizeof(TcItem) * n, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);                     // `items` is on this stack frame
  if (e != cudaSuccess) { set_error("two-tile plan: upload failed: %s", cudaGetErrorString(e)); return NDP_ERR_CUDA; }
  return NDP_OK;
}

void tc_plan_destroy(TcPlan* p) {
  if (!p) return;
  if (p->stream) cudaFree(p->stream);
  if (p->items) cudaFree(p->items);
  if (p->stream2) cudaFree(p->stream2);
  if (p->items2) cudaFree(p->items2);
  delete p;
}

#ifdef NDP_TC_PROF
extern "C" int ndp_debug_tc_prof(unsigned long long* out8, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out8, g_tc_prof, sizeof(unsigned long long) * 8);
  cudaMemcpyFromSymbol(out8 + 8, g_tc_item_wait, sizeof(unsigned long long) * 64);
  cudaMemcpyFromSymbol(out8 + 72, g_tc_a_wait, sizeof(unsigned long long) * 16);
  cudaMemcpyFromSymbol(out8 + 88, g_t2_prof, sizeof(unsigned long long) * 8);
  if (reset) {
    unsigned long long z[64] = {0};
    cudaMemcpyToSymbol(g_tc_prof, z, 64); cudaMemcpyToSymbol(g_tc_item_wait, z, 512); cudaMemcpyToSymbol(g_tc_a_wait, z, 128);
    cudaMemcpyToSymbol(g_t2_prof, z, 64);
  }
  return 0;
}
#endif

// ---- balanced last waves of the two-tile kernel.  P pairs on S SMs (one CTA per SM) take w = ceil(P/S) waves and the
// last one is only partly filled (cfg2: 256 pairs on 148 SMs = 1.73 waves, 16 steps of makespan for 13.8 of work).
// H = w*S - P SMs would idle through it: instead they ("helpers") roll the first `a` steps of m = ceil(n/H) of the
// n = S - H last-wave pairs BEFORE their own last full pair, and each of the other n SMs appends the remaining T - a
// steps of one of those pairs to its own: makespan (w-1)*T + max(T - a, m*a) instead of w*T (cfg2: 14 instead of 16).
struct T2Split { int n, base, h, a; };
static T2Split t2_split(long long P, int T) {             // T <= 0: sizing only (n, base, h)
  T2Split r{0, 0, 0, 0};
  static const bool off = getenv("NDP_TC2_NOSPLIT") != nullptr;
  const long long S = num_sms();
  if (off || S <= 0 || P <= S || P % S == 0) return r;
  const long long w = (P + S - 1) / S, H = w * S - P, n = S - H;
  const long long m = (n + H - 1) / H;
  if ((w - 2) * S > 0x7FFFFFFF) return r;
  long long best = T, best_a = 0;
  for (long long a = 1; a < T; ++a) {
    const long long len = (T - a) > m * a ? (T - a) : m * a;
    if (len < best) { best = len; best_a = a; }
  }
  if (T > 0 && best_a == 0) return r;                      // no split point shortens the last wave (T = 1, or m >= T)
  r.n = (int)n; r.base = (int)((w - 2) * S); r.h = (int)H; r.a = (int)best_a;
  return r;
}
static constexpr size_t kT2SplitFlags = 2048;               // 2 flags per split pair, < 148 pairs
static size_t t2_split_bytes(const TcPlan* p, int G, int K) {
  if (!(p->np == 1 && p->stream2) || G <= 0 || K <= 0) return 0;
  const long long P = (long long)G * (((K + 127) / 128 + 1) / 2);
  const T2Split sp = t2_split(P, 0);                         // n does not depend on T
  return sp.n > 0 ? kT2SplitFlags + (size_t)sp.n * 131072 + 256 : 0;
}
size_t tc_workspace_bytes(const TcPlan* p, int G, int K) {
  return ((size_t)(G > 0 ? G : 1) * p->gext_stride + 255) / 256 * 256 + 256 + t2_split_bytes(p, G, K);
}

template <int NP>
static int launch_np(const TcParams& tp, int grid, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_devs{0};
  if (first_use_on_device(attr_devs)) {
    NDP_CUDA(cudaFuncSetAttribute(rollout_tc_kernel<NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTcSmemBytes));
  }
  rollout_tc_kernel<NP><<<grid, kTcThreads, kTcSmemBytes, s>>>(tp);
  NDP_LAUNCH_CHECK();
  return NDP_OK;
}

int launch_rollout_tc(const TcPlan* P, const float* e0, const float* goal, const float* noise, int G, int K, int T,
                      float* cost, float* a0, float* state_T, void* ws, size_t ws_bytes, cudaStream_t s,
                      const int* list, int R, unsigned long long* packed_c, const int* skip, bool balance) {
  const size_t need = (size_t)G * P->gext_stride;
  if (!ws || ws_bytes < need) { set_error("rollout_score: workspace %zu < %zu", ws_bytes, need); return NDP_ERR_WORKSPACE; }
  NDP_CHECK_ARG(((uintptr_t)ws & 15) == 0, "rollout_score: workspace must be 16-byte aligned");
  // the split hand-over area sits behind the per-goal ext blocks when the caller's workspace has room for it
  // (tc_workspace_bytes); its flags are cleared by the ext kernel, i.e. in stream order before every rollout
  const bool two_tile = P->np == 1 && P->stream2 && !list && !skip;
  const size_t sp_off = (need + 255) / 256 * 256;
  T2Split sp{0, 0, 0, 0};
  if (two_tile && balance) {
    sp = t2_split((long long)G * (((K + 127) / 128 + 1) / 2), T);
    if (sp.n > 0 && ws_bytes < sp_off + kT2SplitFlags + (size_t)sp.n * 131072) sp.n = 0;
  }
  unsigned* sp_flags = sp.n > 0 ? reinterpret_cast<unsigned*>(reinterpret_cast<uint8_t*>(ws) + sp_off) : nullptr;
  tc_goal_ext_kernel<<<dim3(G, 16), 256, 0, s>>>(P->W1, P->b1, goal, 128, P->nd, P->np == 3, reinterpret_cast<__half*>(ws),
                                        P->gext_stride / 2, sp_flags, 2 * sp.n);
  NDP_LAUNCH_CHECK();
  TcParams tp;
  tp.sp_n = sp.n; tp.sp_base = sp.base; tp.sp_h = sp.h; tp.sp_a = sp.a; tp.sp_flags = sp_flags;
  tp.sp_state = sp.n > 0 ? reinterpret_cast<uint8_t*>(ws) + sp_off + kT2SplitFlags : nullptr;
  tp.stream = P->stream; tp.items = P->items; tp.gext = reinterpret_cast<const uint8_t*>(ws);
  tp.gext_stride = P->gext_stride; tp.e0 = e0; tp.goal = goal; tp.noise = noise; tp.cost = cost; tp.a0 = a0;
  tp.state_T = state_T; tp.G = G; tp.K = K; tp.T = T; tp.nd = P->nd; tp.n_items = P->n_items;
  tp.list = list; tp.R = R; tp.packed_c = packed_c; tp.skip = skip;
  tp.tiles_per_goal = ((list ? R : K) + 127) / 128; tp.n_hid_chunks = P->n_chunks;
  if (two_tile) {
    // bulk estimate: two tiles per CTA (rollout_tc2_kernel)
    tp.stream = P->stream2; tp.items = P->items2; tp.n_items = P->n_items2; tp.n_hid_chunks = P->n_chunks2;
    tp.tiles_per_goal = (tp.tiles_per_goal + 1) / 2;            // tile pairs
    const long long grid2 = (long long)G * tp.tiles_per_goal + sp.n;      // a split pair is two CTAs
    NDP_CHECK_ARG(grid2 + 1 <= 0x7FFFFFFF, "rollout_score: G*K too large for one launch");
    tp.n_ctas = (int)grid2;
    { const char* d = getenv("NDP_TC2_DBG"); tp.dbg = d ? atoi(d) : 0; }
    static std::atomic<unsigned long long> attr_devs2{0};
    if (first_use_on_device(attr_devs2)) {
      NDP_CUDA(cudaFuncSetAttribute(rollout_tc2_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kT2SmemBytes));
      NDP_CUDA(cudaFuncSetAttribute(rollout_tc2_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kT2SmemBytes));
      NDP_CUDA(cudaFuncSetAttribute(rollout_tc2_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kT2SmemBytes));
    }
    // clusters of two CTAs share the weight stream (TMA multicast).  The CTAs of a cluster must run the same number of
    // steps (they feed each other's ring), so with a time split its groups (n, H) must be even; the grid is padded to
    // an even size with a ghost CTA
    // MEASURED SLOWER (cfg2: 0.337 vs 0.326 ms; the issuer's wait for ring slots grows from 4.6 k to 5.4 k cycles per
    // pair-step: the wait is latency-, not bandwidth-bound, and a slot now waits for BOTH CTAs) — off unless
    // NDP_TC2_CLUSTER=1; bit-identical results either way (tools/tc2_check.py)
    static const bool want_cluster = getenv("NDP_TC2_CLUSTER") != nullptr;
    const bool cluster = want_cluster && grid2 >= 2 && (sp.n == 0 || (sp.n % 2 == 0 && sp.h % 2 == 0 && sp.base % 2 == 0));
    if (!cluster) {
      if (sp.n > 0) rollout_tc2_kernel<1, true><<<(int)grid2, kT2Threads, kT2SmemBytes, s>>>(tp);
      else rollout_tc2_kernel<1, false><<<(int)grid2, kT2Threads, kT2SmemBytes, s>>>(tp);
    } else {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)((grid2 + 1) / 2 * 2));
      cfg.blockDim = dim3(kT2Threads);
      cfg.dynamicSmemBytes = kT2SmemBytes;
      cfg.stream = s;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      if (tp.dbg & 2) { NDP_CUDA(cudaLaunchKernelEx(&cfg, rollout_tc2_kernel<1, true>, tp)); }   // dev: plain kernel, cluster launch
      else NDP_CUDA(cudaLaunchKernelEx(&cfg, rollout_tc2_kernel<2, true>, tp));
    }
    NDP_LAUNCH_CHECK();
    return NDP_OK;
  }
  const long long grid = (long long)G * tp.tiles_per_goal;
  NDP_CHECK_ARG(grid <= 0x7FFFFFFF, "rollout_score: G*K too large for one launch");
  if (P->np == 1) return launch_np<1>(tp, (int)grid, s);
  if (P->np == 2) return launch_np<2>(tp, (int)grid, s);
  return launch_np<3>(tp, (int)grid, s);
}

}  // namespace ndp
