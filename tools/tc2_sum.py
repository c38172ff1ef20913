"""Checksums of the two-tile K1 kernel's outputs (bulk estimate of precision "f16+refine") on a few shapes — printed so
that two builds can be compared bit for bit on the same box (tools/gpu_run_ab2.sh).  noise_dim 8 / 15 take the wide
ring map (5 x 20 KB slots), the others the narrow one (6 x 18 KB)."""
import hashlib
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from ndivplanning_b200 import FusedPlanner, synthetic

for (G, K, T, H, nd) in [(1, 256, 1, 128, 2), (2, 300, 4, 256, 2), (3, 130, 3, 384, 5), (1, 257, 5, 2048, 15),
                         (2, 1000, 3, 256, 8), (1, 700, 4, 512, 7), (1, 65536, 8, 1024, 2), (1, 40000, 8, 256, 2),
                         (2, 512, 2, 200, 0)]:
    gen, fwd = synthetic.make_models(nd=nd, E=128, H=H, seed=0)
    e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, nd, 128, seed=K + T))
    pl = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine")
    cost, a0, _ = pl.rollout_score(e0, goal, noise)
    torch.cuda.synchronize()
    h = hashlib.sha256(cost.cpu().numpy().tobytes() + a0.cpu().numpy().tobytes()).hexdigest()[:16]
    print(f"sum G={G} K={K} T={T} H={H} nd={nd}: {h} finite={bool(torch.isfinite(cost).all())}", flush=True)
