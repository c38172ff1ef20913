"""Dev: two-tile rollout, cluster vs no-cluster (child processes), tiny set of shapes."""
import os, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from ndivplanning_b200 import FusedPlanner, synthetic
CFGS = [(1, 257, 1, 128, 2), (1, 512, 2, 256, 2), (1, 65536, 8, 1024, 2)]
def run(path):
    out = {}
    for i, (G, K, T, H, nd) in enumerate(CFGS):
        gen, fwd = synthetic.make_models(nd=nd, E=128, H=H)
        e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, nd, 128, seed=5))
        pl = FusedPlanner(gen, fwd, horizon=T, precision="f16")
        c, a, _ = pl.rollout_score(e0, goal, noise)
        c2, a2, _ = pl.rollout_score(e0, goal, noise)
        torch.cuda.synchronize()
        out[i] = (c.cpu(), a.cpu(), bool(torch.equal(c, c2)))
    torch.save(out, path)
if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1]); sys.exit(0)
    with tempfile.TemporaryDirectory() as d:
        res = {}
        for name, env in (("nocluster", {}), ("cluster", dict(NDP_TC2_CLUSTER="1")), ("cluster_localcopy", dict(NDP_TC2_CLUSTER="1", NDP_TC2_DBG="1")), ("plain_kernel_cluster_launch", dict(NDP_TC2_CLUSTER="1", NDP_TC2_DBG="2"))):
            pth = os.path.join(d, name + ".pt")
            subprocess.check_call([sys.executable, os.path.abspath(__file__), pth], env=dict(os.environ, NDP_TC2_NOSPLIT="1", **env), timeout=120)
            res[name] = torch.load(pth)
        for i, cfg in enumerate(CFGS):
            for name in ("cluster", "cluster_localcopy", "plain_kernel_cluster_launch"):
                r, b = res[name][i], res["nocluster"][i]
                bad = (r[0] != b[0])
                print(cfg, name, "equal", bool(torch.equal(r[0], b[0])), "repeatable", r[2], "bad costs", int(bad.sum()),
                      "first bad", bad.flatten().nonzero()[:4].flatten().tolist(), flush=True)
