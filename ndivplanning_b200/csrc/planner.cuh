// ndp_planner handle + device-side parameter block shared by the rollout kernels.
#pragma once
#include "common.cuh"

namespace ndp {

struct RolloutParams {
  const float* wstream;   // fp32 weight stream of one step, 8 KB chunks in consumption order
  const float* b1g;       // [G][128]   gen.fc1 bias + goal half of gen.fc1 (per goal, in workspace)
  const float* w1n;       // [nd][128]  noise columns of gen.fc1, transposed
  const float* b2;        // [64]
  const float* b3;        // [128]
  const float* b4;        // [256]
  const float* w5;        // [4][256]   gen.fc5 (torch layout)
  const float* b5;        // [4]
  const float* wf1a;      // [4][H]     action columns of fwd.fc1, transposed
  const float* bf1;       // [H]
  const float* bf2;       // [E]
  const float* e0;        // [G][E]
  const float* goal;      // [G][E]
  const float* noise;     // [T][G][K][nd]
  float* cost;            // [G][K]
  float* a0;              // [G][K][4]
  float* state_T;         // [G][K][E] or null
  int G, K, T, nd, H, E, tiles_per_goal, nq;   // E: 128 or 256; tiles of rollout_f32_tile(E) candidates
};

int launch_rollout_f32(const RolloutParams& p, cudaStream_t s);
int rollout_f32_smem_bytes(int E);
int rollout_f32_tile(int E);             // candidates per CTA of the fp32 kernel

// ---- tensor-core rollout (k1_rollout_tc.cu) --------------------------------------------------
// One stream item = one weight block landing in one ring slot, consumed by n_main (+1 ext) K=16 slabs.
struct TcItem {
  uint32_t src_off;     // byte offset of the block in the packed weight stream
  uint32_t bytes;       // bytes taken from the stream: main hi [+ main lo] [+ ext hi [+ ext lo]]
  uint32_t ext_bytes;   // bytes taken from the per-goal buffer instead (generator fc1 ext slab), else 0
  uint32_t idesc;       // tcgen05 instruction descriptor (kind::f16, M=128, N)
  uint16_t N;           // GEMM width
  uint16_t d_col;       // TMEM column of the accumulator
  uint16_t a_col;       // TMEM column of the first main slab's hi operand (slabs are 8 columns apart)
  uint16_t lo_delta;    // column offset from a hi slab to its lo slab
  uint8_t n_main;       // main slabs (A operand in TMEM); never straddles an epilogue half
  uint8_t ext_kind;     // 0 none, 1 noise, 2 action, 3 ones
  uint8_t first;        // first item of its GEMM: first MMA overwrites the accumulator
  uint8_t wait_a;       // wait for the next "A ready" event before issuing
  uint8_t signal_d;     // commit the next "D full" event after issuing
  uint8_t pad[3];
};
static_assert(sizeof(TcItem) == 32, "TcItem must stay 32 bytes");
static constexpr int kTcMaxItems = 96;

struct TcParams {
  const uint8_t* stream;    // packed fp16 weight stream of one step
  const TcItem* items;      // [n_items] (device)
  const uint8_t* gext;      // per-goal generator-fc1 ext blocks
  uint32_t gext_stride;     // bytes per goal
  const float* e0;
  const float* goal;
  const float* noise;
  float* cost;
  float* a0;
  float* state_T;
  // list mode (refine pass): candidates of goal g are list[g*R + pos], pos < R (-1 = none); costs/actions
  // are scattered back into cost/a0 and the packed keys (okey(cost)<<32 | candidate) go to packed_c[g*R+pos]
  const int* list;
  unsigned long long* packed_c;
  int R;
  const int* skip;          // nullable [G]: CTAs of goals with skip[g] != 0 exit at once (fallback pass)
  int G, K, T, nd, n_items, tiles_per_goal, n_hid_chunks;
  // two-tile kernel, balanced last waves (k1_rollout_tc.cu: t2_split): sp_n pairs are rolled in two pieces
  // [0, sp_a) and [sp_a, T) by different CTAs; sp_state holds their state images in between, sp_flags[2 * id + tile]
  int sp_n, sp_base, sp_h, sp_a;
  int dbg;                  // development switches (0 in production)
  int n_ctas;               // CTAs that own a job; the grid may be padded to a multiple of the cluster size ("ghost" CTAs
                            // mirror the last job without writing anything, to keep the cluster's weight ring in step)
  unsigned* sp_flags;
  uint8_t* sp_state;
};

struct TcPlan;   // host-side: packed stream + item table of one precision mode
int tc_plan_create(TcPlan** out, int np, int E, int nd, int H, const float* const* Wt, const float* const* bt,
                   cudaStream_t s);
void tc_plan_destroy(TcPlan* p);
size_t tc_workspace_bytes(const TcPlan* p, int G, int K);
int launch_rollout_tc(const TcPlan* plan, const float* e0, const float* goal, const float* noise, int G, int K,
                      int T, float* cost, float* a0, float* state_T, void* ws, size_t ws_bytes, cudaStream_t s,
                      const int* list = nullptr, int R = 0, unsigned long long* packed_c = nullptr,
                      const int* skip = nullptr, bool balance = true);
int select_masked(const float* cost, int G, int K, int topk, long long idx_base, unsigned long long* pm, int* topk_idx,
                  float* topk_cost, void* ws, size_t ws_bytes, const int* skip, cudaStream_t s);
int mark_fallback(int* verified, int G, cudaStream_t s);

int select_packed(const unsigned long long* packed, int G, int n, int topk, long long idx_base,
                  unsigned long long* pm, int* topk_idx, float* topk_cost, cudaStream_t s);
int verify_refine(const float* est_topR, const unsigned long long* packed_c, const float* exact_topk,
                  const unsigned long long* pm, const float* goal, int E, const float* errc, float eps, int G, int K, int R,
                  int topk, int* verified, float* margin, cudaStream_t s);

}  // namespace ndp

struct ndp_planner {
  ndp_planner_desc d;     // d.H is the PADDED hidden width (multiple of 128); H_user is what the caller described
  int H_user;
  bool loaded;
  float* blob;            // one device allocation holding everything below
  size_t blob_floats;
  // offsets (floats) into blob
  float *wstream, *w1g /*[128][E] goal cols of fc1, torch layout rows=n*/, *b1, *w1n, *b2, *b3, *b4, *w5, *b5,
      *wf1a, *bf1, *bf2;
  // plain torch-layout copies for the single-call APIs (generator_forward / forward_step)
  float* Wt[7];
  float* bt[7];
  float* scratch;         // activations for the single-call APIs, grown on demand
  size_t scratch_floats;
  int nq;
  float* errc;            // device scalar: a-priori rounding floor of the single-pass estimate in sqrt-cost space, per unit
                          // of state RMS (computed from the forward model's row norms at load; see verify_refine_kernel)
  ndp::TcPlan* tc;     // tensor-core plan (precision != fp32), else null
  ndp::TcPlan* tc_refine;  // f16x3 plan of the refine pass (NDP_PREC_F16_REFINE), else null
};
