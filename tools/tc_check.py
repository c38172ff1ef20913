"""Dev check: tensor-core rollout modes against the fp32 CUDA mode (itself parity-tested vs the oracle)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ndivplanning_b200 import FusedPlanner, synthetic

modes = sys.argv[1].split(",") if len(sys.argv) > 1 else ["f16", "f16x2", "f16x3"]
dev = torch.device("cuda", 0)
CFGS = [(1, 1024, 8, 1024), (2, 300, 3, 256), (1, 65536, 8, 1024)]
if len(sys.argv) > 2:
    CFGS = [CFGS[int(sys.argv[2])]]
for (G, K, T, H) in CFGS:
    gen, fwd = synthetic.make_models(nd=2, E=128, H=H)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, 2, 128, seed=5)
    e0, goal, noise = e0.to(dev), goal.to(dev), noise.to(dev)
    ref = FusedPlanner(gen, fwd, horizon=T, precision="fp32", device=dev)
    c_ref, a_ref, s_ref = ref.rollout_score(e0, goal, noise, want_state=True)
    torch.cuda.synchronize()
    for mode in modes:
        pl = FusedPlanner(gen, fwd, horizon=T, precision=mode, device=dev)
        c, a, s = pl.rollout_score(e0, goal, noise, want_state=True)
        torch.cuda.synchronize()
        rel = ((c - c_ref).abs() / c_ref.abs()).max().item()
        arel = ((a - a_ref).abs().max() / a_ref.abs().max()).item()
        srel = ((s - s_ref).abs().max() / s_ref.abs().max()).item()
        same = bool((c.argmin(1) == c_ref.argmin(1)).all())
        ts = []
        for _ in range(5):
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record(); pl.rollout_score(e0, goal, noise); b_.record(); torch.cuda.synchronize()
            ts.append(a_.elapsed_time(b_))
        ms = min(ts)
        tp = []
        for _ in range(5):
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record(); pl.plan(e0, goal, noise, topk=16); b_.record(); torch.cuda.synchronize()
            tp.append(a_.elapsed_time(b_))
        print(f"G={G} K={K} T={T} H={H} {mode:6s} cost maxrel={rel:.3e} a0 rel={arel:.3e} state rel={srel:.3e} "
              f"argmin same={same} finite={bool(torch.isfinite(c).all())}  {ms:.3f} ms  {G*K*T/ms/1e6:.2f} Gsteps/s  plan(topk16) {min(tp):.3f} ms", flush=True)
