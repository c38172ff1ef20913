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
  int G, K, T, nd, H, tiles_per_goal, nq;
};

int launch_rollout_f32(const RolloutParams& p, cudaStream_t s);
int rollout_f32_smem_bytes();

}  // namespace ndp

struct ndp_planner {
  ndp_planner_desc d;
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
};
