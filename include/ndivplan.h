/* libndivplan — C ABI of the B200 (sm_100a) planning hot path of ndivplanning.
 *
 * The reference (goodmattg/ndivplanning) is pure Python/PyTorch and has no FFI
 * layer; each entry point below names the reference lines it replaces.  Paths
 * are relative to the reference tree.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - all tensors fp32, row-major contiguous; indices int32; packed keys uint64;
 *   - stream is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - all work is stream-ordered, no host synchronisation inside a call;
 *   - return 0 on success, <0 on error; text via ndp_last_error() (thread-local);
 *   - the caller owns every buffer; the library allocates only inside handles;
 *   - there is NO CPU code path: calls fail with NDP_ERR_DEVICE off sm_100.
 */
#ifndef NDIVPLAN_H_
#define NDIVPLAN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NDP_ABI_VERSION 1

enum {
  NDP_OK = 0,
  NDP_ERR_ARG = -1,      /* bad shape / null pointer / unsupported size */
  NDP_ERR_CUDA = -2,     /* a CUDA runtime call failed */
  NDP_ERR_DEVICE = -3,   /* not an sm_100 device */
  NDP_ERR_WORKSPACE = -4,/* workspace too small */
  NDP_ERR_NCCL = -5
};

/* arithmetic mode of the rollout kernel */
enum {
  NDP_PREC_FP32 = 0,     /* CUDA-core FFMA, fp32 everywhere: parity mode            */
  NDP_PREC_F16X3 = 1,    /* tcgen05 kind::f16, 3-term hi/lo split operands, fp32 acc */
  NDP_PREC_F16 = 2,      /* tcgen05 kind::f16 single pass: bulk estimate (cost ~2e-4 rel) */
  NDP_PREC_F16X2 = 3,    /* tcgen05 kind::f16, activations split hi/lo, weights fp16: cost ~1.5e-4 rel */
  NDP_PREC_F16_REFINE = 4 /* ndp_plan_select only: F16 bulk estimate of all K, then the 128 best estimates are
                             re-rolled in F16X3 and the plan is selected from those fp32-grade costs */
};

typedef struct ndp_planner ndp_planner;
typedef struct ndp_comm ndp_comm;

typedef struct {
  int E;          /* state-embedding width: 128 (image_autoencoder.py:27) or 256 (fp32 mode only) */
  int nd;         /* noise_dim (evaluation.noise_dim; mpc_eval.py:65)                */
  int A;          /* action width, must be 4 (gan.py:71)                             */
  int H;          /* latent forward model hidden width, any value in [1,8192] (padded inside
                     the library to a multiple of 128 with zero units: exact)          */
  int T;          /* mpc.time_horizon (mpc_eval.py:68)                               */
  int precision;  /* NDP_PREC_*                                                      */
} ndp_planner_desc;

int ndp_abi_version(void);
const char* ndp_last_error(void);
/* 0 when the current device is sm_100 and the kernels can run on it */
int ndp_device_check(void);

/* ---- planner: generator MLP + T-step latent rollout + terminal cost -------------
 * replaces mpc_eval.py:124-157 (rollout), :161 (per-candidate MSE) with
 * models/gan.py:79-86 (Decoder.forward) and forward_encoder.py:105-114 semantics. */
int ndp_planner_create(ndp_planner** out, const ndp_planner_desc* desc);
void ndp_planner_destroy(ndp_planner* p);

/* Weights in torch nn.Linear layout [out,in] (device pointers), order:
 *   W[0..4] = generator fc1..fc5  (fc1: [128, 2E+nd], columns = [state | goal | noise],
 *                                  gan.py:67 + cat order mpc_eval.py:137-138,:82)
 *   W[5]    = forward fc1 [H, E+A] (columns = [state | action], forward_encoder.py:107)
 *   W[6]    = forward fc2 [E, H]
 * b[i] the matching biases. Copies + prepacks on `stream`; sources may be freed after
 * the stream reaches this point. */
int ndp_planner_load(ndp_planner* p, const float* const* W, const float* const* b,
                     int n_tensors, void* stream);

int ndp_planner_workspace_bytes(const ndp_planner* p, int G, int K, size_t* bytes);

/* e0,goal [G,E]; noise [T,G,K,nd] (step-major draw order of mpc_eval.py:80,:141);
 * cost [G,K] = mean_E (s_T - goal)^2 (mpc_eval.py:161); a0 [G,K,4] = step-0 actions
 * (mpc_eval.py:150-151); state_T [G,K,E] optional (NULL to skip). */
int ndp_rollout_score(ndp_planner* p, const float* e0, const float* goal, const float* noise,
                      int G, int K, float* cost, float* a0, float* state_T,
                      void* ws, size_t ws_bytes, void* stream);

/* one generator forward (models/gan.py:79-86): z [B, 2E+nd] -> a [B,4] */
int ndp_generator_forward(ndp_planner* p, const float* z, int B, float* a, void* stream);
/* one forward-model step (forward_encoder.py:105-114, eval mode: state + residual) */
int ndp_forward_step(ndp_planner* p, const float* state, const float* action, int B,
                     float* next_state, void* stream);

/* ---- sampler: replaces torch.FloatTensor(N,S,nd).uniform_() of diverse_sampling (mpc_eval.py:78-83) on the device.
 * out[i] = (philox4x32_10(counter = {i/4, offset}, key = seed)[i & 3] >> 8) * 2^-24 in [0,1); reproducible from
 * (seed, offset), independent of launch geometry.  Fill noise [T,G,K,nd] with it and pass it to the plan calls. */
int ndp_sample_uniform(uint64_t seed, uint64_t offset, float* out, size_t n, void* stream);

/* ---- selection: replaces the python scan mpc_eval.py:122,159-165 -----------------
 * cost [G,K]. For each segment g:
 *   packed_min[g] = (okey(cost)<<32) | (idx_base + argmin)  with first-index tie break;
 *                   if no cost is < 1e10 (incl. NaN) -> 0xFFFFFFFF00000000 (index 0, the
 *                   reference's min_error=1e10 / best_act_ind=0 initialisation);
 *   topk_idx[g,:]  = the topk smallest under the total order (cost, index), ascending
 *                   (NaN last) == torch.sort(stable=True)[:topk]; topk<=128; may be NULL
 *                   with topk==0; topk_cost likewise.
 * okey = fp32 bits with the sign bit flipped (non-negative) or all bits flipped (negative).
 * ws: ndp_select_workspace_bytes(). */
int ndp_select_workspace_bytes(int G, int K, int topk, size_t* bytes);
int ndp_select(const float* cost, int G, int K, int topk, int64_t idx_base,
               uint64_t* packed_min, int32_t* topk_idx, float* topk_cost,
               void* ws, size_t ws_bytes, void* stream);
/* ---- ndp_select without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, results and kernels as ndp_select, but the scratch is a file-scope `static __device__` array of
 * ndp_select_static_max_bytes() = 1 MiB inside the library (held by every GPU that loads it, until exit); the entry
 * point calls no cudaMalloc / cudaFree / allocating routine.  Shapes whose ndp_select_workspace_bytes() exceeds the
 * array are rejected with NDP_ERR_WORKSPACE (e.g. G=4096, K=65,536, topk=16; every BASELINE config fits).  All calls on
 * one GPU share the array: what must start at zero is cleared in stream order by the call itself, and calls that
 * OVERLAP on the GPU (different streams) race — serialise them or use ndp_select with one workspace per stream. */
int ndp_select_static_max_bytes(size_t* bytes);
int ndp_select_static(const float* cost, int G, int K, int topk, int64_t idx_base, uint64_t* packed_min,
                      int32_t* topk_idx, float* topk_cost, void* stream);
/* ---- one-call plan: rollout + selection (mpc_eval.py:124-165 minus the action gather) --------------
 * = ndp_rollout_score followed by ndp_select for every precision; with NDP_PREC_F16_REFINE the bulk
 * costs are single-pass estimates, the NDP_REFINE_R best estimates per goal are re-rolled with split
 * operands (their cost / a0 entries are overwritten with the refined values) and the selection is taken
 * from the refined costs.  verified[g] (nullable) = 1 when no candidate outside the refine list can belong to the
 * top-k under the error model of the estimate: |sqrt(estimate) - sqrt(exact)| <= eta for every candidate (an absolute
 * bound on the terminal-state error, valid however close a candidate comes to the goal), with
 *   eta = max( 3 x the largest |sqrt(estimate) - sqrt(refined)| observed on the NDP_REFINE_R re-rolled candidates of
 *              this plan,  an a-priori first-order rounding floor from the forward model's row norms and T,
 *              0.5 * eps * sqrt(estimate_R) when eps > 0 )
 * and the test sqrt(k-th refined cost) < sqrt(R-th estimate) - eta.  eps <= 0: no caller floor (the default); eps > 0
 * can only make the test stricter.  Always 1 for the non-refined precisions.  topk <= 64 in refine mode.
 * The bound is derived from the data of the plan, not proven a priori: goals that fail it are NOT silently accepted
 * (verified = 0) and ndp_plan_fallback re-rolls them exactly.
 * ndp_plan_fallback (same arguments, call after ndp_plan_select / _refine on the same stream): goals with
 * verified[g] == 0 are re-rolled over ALL K candidates with split operands and selected from those costs;
 * the decision is a device-side mask (no host round trip; thread blocks of verified goals exit at once) and
 * such goals then report verified[g] = 2.  No-op for the non-refined precisions. */
#define NDP_REFINE_R 128
int ndp_plan_workspace_bytes(const ndp_planner* p, int G, int K, int topk, size_t* bytes);
int ndp_plan_select(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                    int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                    int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                    void* stream);
/* The same call split in two stream-ordered halves with identical arguments, so that a caller may overlap the
 * (latency-bound, single-tile) refine half of plan i with the bulk half of plan i+1 on another stream:
 *   _bulk  : rollout of all K (+ the whole selection when not refining);
 *   _refine: selection of the refine list, re-roll, selection from refined costs, verification (no-op when
 *            not refining).
 * The caller orders _refine after _bulk (event) and keeps the workspace alive across both. */
int ndp_plan_fallback(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                      int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                      int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                      void* stream);
int ndp_plan_select_bulk(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                         int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                         int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                         void* stream);
int ndp_plan_select_refine(ndp_planner* p, const float* e0, const float* goal, const float* noise, int G, int K,
                           int topk, int64_t idx_base, float eps, float* cost, float* a0, uint64_t* packed_min,
                           int32_t* topk_idx, float* topk_cost, int32_t* verified, void* ws, size_t ws_bytes,
                           void* stream);

/* after ndp_plan_select / _refine on `stream` with the same (G, K, topk, ws): info[g] = {margin, eta} in sqrt-cost
 * units — margin = sqrt(R-th estimate) - eta - sqrt(k-th refined cost) (> 0 <=> verified), eta as defined above */
int ndp_plan_verify_info(const ndp_planner* p, int G, int K, int topk, const void* ws, size_t ws_bytes, float* info,
                         void* stream);

/* best_action[g,:] = a0[g, idx(packed_min[g]) - idx_base, :]   (mpc_eval.py:165) */
int ndp_gather_best_action(const float* a0, const uint64_t* packed_min, int G, int K,
                           int64_t idx_base, float* best_action, int32_t* best_idx,
                           void* stream);

/* ---- sharded plan exchange (replaces nothing in the reference, which is single-device) ----
 * A rank's local plan is packed into one record of (4 + topk) u64 words per goal:
 *   [0] packed_min, [1..2] best_action (4 x f32 bits), [3] verified, [4..) top-k keys
 *   (okey(cost)<<32 | GLOBAL index; missing entries all ones).
 * The caller all-gathers the records (ONE collective per plan) into rec_all [P,G,4+topk] and
 * ndp_shard_merge reduces them: unsigned min of [0] with lowest-rank tie break (ranks own increasing
 * index ranges, so this is the reference's first-index rule, mpc_eval.py:159-164), the owner's action,
 * AND of verified, and the global top-k under (cost, index).  P*topk <= 1024. */
int ndp_shard_pack(const uint64_t* packed_min, const float* best_action, const int32_t* verified,
                   const int32_t* topk_idx, const float* topk_cost, int G, int topk, uint64_t* rec,
                   void* stream);
int ndp_shard_merge(const uint64_t* rec_all, int P, int G, int topk, uint64_t* packed_min, int32_t* best_idx,
                    float* best_action, int32_t* topk_idx, float* topk_cost, int32_t* verified, void* stream);
/* pack -> exchange -> merge as ONE stream-ordered call on communicator `c` (a collective: every rank calls it once per
 * plan, in the same order).  Over the peer window (ndp_comm_peer_status) it is two launches and no library collective:
 * the pack kernel stores the record words straight into EVERY rank's window over NVLink and publishes a flag, the merge
 * kernel waits (bounded) for every rank's flag and merges from its own window; up to 8 exchanges may be in flight
 * (PlanRing depth <= 7).  Otherwise: ndp_shard_pack -> ncclAllGather -> ndp_shard_merge through the scratch buffers
 * rec [G,4+topk] and rec_all [world,G,4+topk] (may be NULL only if the window path is known to be active). */
int ndp_shard_exchange(ndp_comm* c, const uint64_t* packed_min, const float* best_action, const int32_t* verified,
                       const int32_t* topk_idx, const float* topk_cost, int G, int topk, uint64_t* rec,
                       uint64_t* rec_all, uint64_t* out_packed_min, int32_t* out_best_idx, float* out_best_action,
                       int32_t* out_topk_idx, float* out_topk_cost, int32_t* out_verified, void* stream);

/* ---- normalized diversification loss: replaces diversity.py:8-41 -----------------
 * x = recodes [N,k,Dx] (outputs), z = codes [N,k,Dz] (latents) — argument order of
 * compute_pairwise_divergence(recodes, codes), diversity.py:36.
 * loss[0] = sum relu(margin * dz_ij/sz_i - dx_ij/sx_i); sx,sz [N,k] row sums (detached
 * normalisers, diversity.py:17-18). grad_x (nullable) = d loss / d x for grad_loss = 1.
 * Row-block sharding (multi-GPU): rows [row0,row0+rows) of every group are owned by the
 * caller; sx/sz must hold ALL rows before ndp_ndiv_loss_grad (allgather in between).
 * ws: ndp_ndiv_workspace_bytes() bytes, used by both passes (partial sums of the column splits). */
int ndp_ndiv_workspace_bytes(int N, int k, int Dx, int Dz, size_t* bytes);
int ndp_ndiv_rowsums(const float* x, const float* z, int N, int k, int Dx, int Dz,
                     int row0, int rows, float* sx, float* sz, void* ws, size_t ws_bytes, void* stream);
int ndp_ndiv_loss_grad(const float* x, const float* z, const float* sx, const float* sz,
                       int N, int k, int Dx, int Dz, float margin, int row0, int rows,
                       float* loss, float* grad_x, void* ws, size_t ws_bytes, void* stream);
/* both passes over all rows in one call (single-GPU autograd step; operands are packed once):
 * sx, sz [N,k] out, loss[0], grad_x nullable */
int ndp_ndiv_forward_backward(const float* x, const float* z, int N, int k, int Dx, int Dz, float margin,
                              float* sx, float* sz, float* loss, float* grad_x, void* ws, size_t ws_bytes,
                              void* stream);
/* the operator as a forward / backward pair (what a torch.autograd.Function binds):
 * forward  -> loss[0] + the saved normalisers sx, sz [N,k];
 * backward -> grad_x (nullable) = grad_loss[0] * d loss / d x (grad_loss: device scalar, NULL = 1); grad_z (nullable) =
 *             grad_loss[0] * d loss / d z, i.e. what autograd gives when `codes.requires_grad` (diversity.py:36-41; the
 *             reference's own call site passes raw noise and never asks for it, train_gan.py:193-196) — exact direct
 *             differences through an [N,k,k] scratch: ws must then hold ndp_ndiv_gradz_workspace_bytes();
 *             loss_scratch: one device float */
int ndp_ndiv_gradz_workspace_bytes(int N, int k, int Dx, int Dz, size_t* bytes);
int ndp_ndiv_forward(const float* x, const float* z, int N, int k, int Dx, int Dz, float margin, float* loss,
                     float* sx, float* sz, void* ws, size_t ws_bytes, void* stream);
int ndp_ndiv_backward(const float* x, const float* z, const float* sx, const float* sz, int N, int k, int Dx, int Dz,
                      float margin, const float* grad_loss, float* grad_x, float* grad_z, float* loss_scratch,
                      void* ws, size_t ws_bytes, void* stream);
/* compute_pairwise (normalise=0, diversity.py:8-9) / compute_pair_distance
 * (normalise=1, diversity.py:12-19): v [N,k,C] -> out [N,k,k] */
int ndp_pairwise(const float* v, int N, int k, int C, int normalise, float* out, void* stream);
/* grad_v [N,k,C] of sum(out * grad_out) for compute_pairwise / compute_pair_distance;
 * rowsum_ws: [N,k] scratch, required when normalise=1 */
int ndp_pairwise_backward(const float* v, const float* grad_out, int N, int k, int C,
                          int normalise, float* rowsum_ws, float* grad_v, void* stream);

/* ---- GAN training step around the ndiv loss -----------------
 * One ndp_gan_train_step == one iteration of the loop body train_gan.py:159-203 on encoded inputs:
 *   a_hat = G([codes | noise])  (Decoder, gan.py:63-86);  d_steps x { D_loss = BCE(D(actions, codes), 1) +
 *   BCE(D(a_hat, codes), 0); Adam(D) }  (Discriminator, gan.py:89-110; train_gan.py:172-184);  G_loss =
 *   BCE(D(a_hat, codes), 1) (:187-190);  div = compute_pairwise_divergence(a_hat[F,S,A], noise[F,S,nd]) (:193-196);
 *   Adam(G) on G_loss + div_factor * div (:198-203).  Optimisers: torch.optim.Adam(lr, betas) of :98-104 (the
 *   encoder parameters in G_optimizer never receive a gradient: its outputs are detached, :152-153).
 * The handle owns parameters, Adam state and all scratch (sized for max_F codes); all work is stream-ordered
 * without host synchronisation (capturable in a CUDA graph).  The layer widths are the reference's:
 * G: C+nd -> 128 -> 64 -> 128 -> 256 -> A (ReLU), D: A+C -> 64 -> 128 -> 256 -> 1 (leaky ReLU 0.01). */
typedef struct ndp_gan ndp_gan;
typedef struct {
  int32_t C;        /* conditioning width (256 = state | goal embedding) */
  int32_t nd;       /* noise_dim */
  int32_t A;        /* action dim (4) */
  int32_t S;        /* num_sample: samples per code, the ndiv group size k */
  int32_t max_F;    /* largest number of codes per step (batch_size * (seq_length-1)) */
  int32_t d_steps;  /* discrim_steps_per_gen */
  float margin;     /* 0.8 (diversity.py:40) */
  float div_factor; /* pairwise_div_factor */
  double lr, beta1, beta2, eps;   /* Adam: 2e-4, 0.5, 0.999, 1e-8 */
} ndp_gan_desc;
int ndp_gan_create(ndp_gan** out, const ndp_gan_desc* d);
void ndp_gan_destroy(ndp_gan* g);
/* device pointers to torch [out,in] weights / biases: 5 generator layers, 4 discriminator layers; copies them
 * and resets both optimisers (step 0, zero moments) */
int ndp_gan_load(ndp_gan* g, const float* const* gW, const float* const* gb, const float* const* dW,
                 const float* const* db, void* stream);
int ndp_gan_export(ndp_gan* g, float* const* gW, float* const* gb, float* const* dW, float* const* db,
                   void* stream);
/* codes [F,C], actions [F,A], noise [F,S,nd]; losses[3] = {D_loss, G_loss, pair_div_loss} (all device fp32) */
int ndp_gan_train_step(ndp_gan* g, const float* codes, const float* actions, const float* noise, int F,
                       float* losses, void* stream);
/* data-parallel variant (SURVEY.md 8e for this row): same parameters on every rank, F codes of a world*F global
 * batch per rank; ONE all-reduce of the flat gradient per optimiser step + one of the 3 loss scalars, on `stream` */
int ndp_gan_train_step_dp(ndp_gan* g, ndp_comm* comm, int world, const float* codes, const float* actions,
                          const float* noise, int F, float* losses, void* stream);
int ndp_gan_set_lr(ndp_gan* g, double lr, void* stream);        /* StepLR (train_gan.py:105) lives in the caller */
/* inspection: flat [W0|b0|W1|b1|...] gradients of the last step (either may be NULL), a_hat [F*S,A] */
int ndp_gan_param_counts(const ndp_gan* g, size_t* n_gen, size_t* n_dis);
int ndp_gan_grads(ndp_gan* g, float* gen_flat, float* dis_flat, void* stream);
int ndp_gan_last_actions(ndp_gan* g, float* a_hat, int F, void* stream);

/* ---- image-encoder front end: models/image_autoencoder.py:14-49 (Encoder) in eval mode,
 * [B,3,128,128] NCHW frames -> [B,128] codes (the reference's [B,128,1,1]); replaces image_encoder(x) at
 * mpc_eval.py:135-136 / train_gan.py:152-153.  ndp_encoder_load takes, per layer, weights permuted to
 * [Cout][KH][KW][Cin] with the eval-mode BatchNorm of conv1..conv3 folded in, and the folded biases. */
typedef struct ndp_encoder ndp_encoder;
int ndp_encoder_create(ndp_encoder** out, int max_images, int img /* 128 */);
void ndp_encoder_destroy(ndp_encoder* e);
int ndp_encoder_load(ndp_encoder* e, const float* const* W /*[6]*/, const float* const* b /*[6]*/, void* stream);
int ndp_encoder_forward(ndp_encoder* e, const float* images, int B, float* codes, void* stream);

/* ---- collectives (NCCL over NVLink; libnccl is dlopen'ed on first use) ----------- */
int ndp_comm_unique_id(void* id128_host);                       /* 128-byte ncclUniqueId */
int ndp_comm_create(ndp_comm** out, const void* id128_host, int rank, int world);
int ndp_comm_allreduce_min_u64(ndp_comm* c, uint64_t* buf, int count, void* stream);
int ndp_comm_allreduce_sum_f32(ndp_comm* c, float* buf, int count, void* stream);
int ndp_comm_allgather_f32(ndp_comm* c, const float* send, float* recv, int count_per_rank,
                           void* stream);
/* rank-major all-gather of opaque bytes (the sharded plan's record exchange) */
int ndp_comm_allgather_bytes(ndp_comm* c, const void* send, void* recv, size_t bytes_per_rank,
                             void* stream);
int ndp_comm_rank(const ndp_comm* c);
int ndp_comm_world(const ndp_comm* c);
/* Peer window (csrc/peer.cuh): on one node every rank maps every other rank's exchange window through CUDA IPC on first
 * use, and the sharded entry points below exchange their results with loads / stores over NVLink and one flag word
 * per (phase, rank) instead of library collectives; ranks that cannot map each other (or NDP_NO_PEER_WINDOW=1, or a
 * stream under graph capture) use the NCCL collectives above.  *mapped = 1 when the window path is active.  Returns
 * NDP_ERR_NCCL if a bounded in-kernel wait for a peer expired since the last call (a rank missed a collective
 * call); synchronises `stream`. */
int ndp_comm_peer_status(ndp_comm* c, int* mapped, void* stream);
void ndp_comm_destroy(ndp_comm* c);

/* ---- row-sharded ndiv loss, forward + backward, one group (diversity.py:12-19,36-41): x [k,Dx], z
 * [k,Dz] replicated on every rank, rank r owns rows [r*B, (r+1)*B), B = ceil(k/world) rounded up to 128.  ONE stream-
 * ordered call per rank: pre-pack, row sums of the local rows, exchange of the [sx|sz] blocks, loss + gradient of
 * the local rows, exchange of [grad rows | loss partial] (peer window: the pass kernels' last CTA publishes the rank's
 * slot, a pull kernel gathers every rank's slot over NVLink — 5 launches, no library collective; otherwise two
 * in-place ncclAllGather); sx, sz [k], loss[0] (sum of the partials in rank order) and
 * grad_x [k,Dx] (nullable) are complete and bit-identical on every rank.  Shapes of the tensor-core path only
 * (k >= 1536, Dx,Dz <= 32, Dx % 4 == 0); ws: ndp_ndiv_sharded_workspace_bytes(), 256-byte aligned. */
int ndp_ndiv_sharded_workspace_bytes(int k, int Dx, int Dz, int world, size_t* bytes);
int ndp_ndiv_sharded_forward_backward(ndp_comm* c, const float* x, const float* z, int k, int Dx, int Dz, float margin,
                                      float* sx, float* sz, float* loss, float* grad_x, void* ws, size_t ws_bytes,
                                      void* stream);

/* number of kernels this library has launched in this process (bench gpu_launches) */
uint64_t ndp_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* NDIVPLAN_H_ */
