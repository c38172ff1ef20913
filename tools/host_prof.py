import sys, os, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ndivplanning_b200 import FusedPlanner, synthetic
dev = torch.device("cuda", 0)
G, K, T, H = 1, 65536, 8, 1024
gen, fwd = synthetic.make_models(nd=2, E=128, H=H)
e0, goal, noise = [x.to(dev) for x in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=5)]
pl = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine", device=dev)
for _ in range(3): pl.plan(e0, goal, noise, topk=16)
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
ps = [pl.plan(e0, goal, noise, topk=16) for _ in range(30)]
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(8)
