"""Secondary row of BASELINE cfg2 under the alternative reading "embedding dim 256 = E" (SURVEY.md 8a note): the fp32
kernel at E = 256 (64-candidate tiles) next to the same kernel at E = 128, cfg2 shape (G=1, K=65,536, T=8, H=512, nd=2).
Prints one JSON line; device time by CUDA events over `reps` back-to-back rollouts after warm-up."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch

from ndivplanning_b200 import FusedPlanner, synthetic

G, K, T, H, nd = 1, 65536, 8, 512, 2
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
out = {"workload": f"G={G} K={K} T={T} H={H} nd={nd}", "precision": "fp32"}
for E in (128, 256):
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    pl = FusedPlanner(gen, fwd, horizon=T, precision="fp32")
    e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, nd, E, seed=5))
    for _ in range(2):
        pl.rollout_score(e0, goal, noise)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(reps):
        pl.rollout_score(e0, goal, noise)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    flop = 2.0 * K * T * ((2 * E + nd) * 128 + 128 * 64 + 64 * 128 + 128 * 256 + 256 * 4 + (E + 4) * H + H * E)
    out[f"E{E}"] = {"ms": round(ms, 4), "cand_steps_per_s": G * K * T / ms * 1e3, "tflops_fp32": flop / ms * 1e-9}
print(json.dumps(out))
