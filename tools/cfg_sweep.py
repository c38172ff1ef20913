"""BASELINE configs 4 and 5 at their per-GPU shard sizes on ONE GPU (weak-scaling units of the 8-GPU runs):
cfg4: K = 1,048,576 / 8 = 131,072 candidates, T = 16, one goal (PlanRing loop);
cfg5: G = 4,096 / 8 = 512 goals x K = 4,096 x T = 8, 7 receding-horizon outer steps (ClosedLoopSweep)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import FusedPlanner, synthetic  # noqa: E402
from ndivplanning_b200.closed_loop import ClosedLoopSweep  # noqa: E402


def ev():
    return torch.cuda.Event(enable_timing=True)


def main():
    dev = torch.device("cuda", 0)
    gen, fwd = synthetic.make_models(nd=2, E=128, H=1024)
    # ---- cfg4 shard
    G, K, T, topk = 1, 131072, 16, 16
    pl = FusedPlanner(gen, fwd, horizon=T, device=dev, precision="f16+refine")
    ring = pl.ring(G, K, topk)
    e0, goal, _ = synthetic.make_plan_inputs(G, 1, 1, 2, 128, seed=1)
    e0, goal = e0.to(dev), goal.to(dev)
    pool = [torch.rand(T, G, K, 2, device=dev) for _ in range(10)]      # 10 x 16.8 MB > L2
    for i in range(4):
        ring.plan(e0, goal, pool[i])
    torch.cuda.synchronize()
    a, b = ev(), ev()
    n = 20
    a.record()
    plans = []
    for i in range(n):
        plans.append(ring.plan(e0, goal, pool[i % 10]))
        if len(plans) > 2:
            plans.pop(0).wait()
    for p in plans:
        p.wait()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    exact = FusedPlanner(gen, fwd, horizon=T, device=dev, precision="fp32").plan(e0, goal, pool[(n - 1) % 10], topk=topk)
    last = plans[-1]
    print(f"cfg4 shard (K={K}, T={T}): {ms:.3f} ms/plan = {G * K * T / (ms * 1e-3):.3e} candidate-steps/s; "
          f"argmin == fp32 path: {bool(torch.equal(last.best_idx, exact.best_idx))}, "
          f"top-{topk} set equal: {set(last.topk_idx[0].tolist()) == set(exact.topk_idx[0].tolist())}, verified {last.verified.tolist()}",
          flush=True)
    del ring, pool, plans, exact
    # ---- cfg5 shard
    G, K, T, S = 512, 4096, 8, 7
    sw = ClosedLoopSweep(gen, fwd, horizon=T, rollouts=K, precision="f16+refine", device=dev)
    torch.manual_seed(3)
    start, goal = torch.randn(G, 128, device=dev), torch.randn(G, 128, device=dev)
    fut = torch.randn(S, G, 128, device=dev)
    gen_d = torch.Generator(device=dev).manual_seed(7)
    sw.run(start, goal, S, state_fut=fut, generator=gen_d)
    torch.cuda.synchronize()
    a, b = ev(), ev()
    a.record()
    res = sw.run(start, goal, S, state_fut=fut, generator=gen_d)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    print(f"cfg5 shard (G={G}, K={K}, T={T}, {S} receding outer steps): {ms:.2f} ms per sweep = "
          f"{res.candidate_steps / (ms * 1e-3):.3e} candidate-steps/s; verified {int(res.verified.sum())}/{res.verified.numel()}",
          flush=True)
    # non-receding (full horizon every step, MPC_gym_eval.py:188)
    a.record()
    res = sw.run(start, goal, S, state_fut=fut, generator=gen_d, receding=False)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    print(f"cfg5 shard, full horizon every step: {ms:.2f} ms per sweep = {res.candidate_steps / (ms * 1e-3):.3e} "
          f"candidate-steps/s", flush=True)


if __name__ == "__main__":
    main()
