"""closed-loop sweep (N1): eager stream-ordered loop vs the whole loop as one CUDA graph.  python tools/sweep_graph_time.py"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import synthetic
from ndivplanning_b200.closed_loop import ClosedLoopSweep

dev = torch.device("cuda", 0)
for G, K, Th, S in ((1, 4096, 8, 7), (16, 4096, 8, 7), (512, 4096, 8, 7)):
    gen, fwd = synthetic.make_models(nd=2, E=128, H=1024)
    torch.manual_seed(0)
    start, goal = torch.randn(G, 128, device=dev), torch.randn(G, 128, device=dev)
    sw = ClosedLoopSweep(gen, fwd, horizon=Th, rollouts=K, precision="f16+refine")
    res = sw.run(start, goal, S); torch.cuda.synchronize()
    n = 10 if G < 512 else 3
    t0 = time.perf_counter()
    for _ in range(n):
        res = sw.run(start, goal, S)
    torch.cuda.synchronize()
    eager = (time.perf_counter() - t0) / n
    g, r = sw.capture(start, goal, S)
    g.replay(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        g.replay()
    torch.cuda.synchronize()
    graph = (time.perf_counter() - t0) / n
    print(f"G={G} K={K} Th={Th} steps={S}: eager {eager*1e3:.3f} ms, one graph {graph*1e3:.3f} ms "
          f"({res.candidate_steps/graph:.3e} candidate-steps/s), verified>=1: {bool((r.verified >= 1).all())}", flush=True)
