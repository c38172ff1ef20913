"""Dev: per-role cycle counters of the tensor-core rollout (library built with -DNDP_TC_PROF)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ndivplanning_b200 import FusedPlanner, synthetic, _lib
dev = torch.device("cuda", 0)
G, K, T, H = 1, 65536, 8, 1024
gen, fwd = synthetic.make_models(nd=2, E=128, H=H)
e0, goal, noise = [x.to(dev) for x in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=5)]
l = _lib.lib()
out = (C.c_ulonglong * 88)()
for mode in sys.argv[1].split(","):
    pl = FusedPlanner(gen, fwd, horizon=T, precision=mode, device=dev)
    pl.rollout_score(e0, goal, noise); torch.cuda.synchronize()
    C.CDLL(_lib.LIB_PATH).ndp_debug_tc_prof(out, 1)
    pl.rollout_score(e0, goal, noise); torch.cuda.synchronize()
    C.CDLL(_lib.LIB_PATH).ndp_debug_tc_prof(out, 1)
    n = 512 * T  # tile-steps
    v = [x / n for x in out]
    print("  wait_full per item:", " ".join(f"{x:.0f}" for x in v[8:8 + 41]))
    print("  wait_a per event  :", " ".join(f"{x:.0f}" for x in v[72:72 + 14]))
    print(f"{mode}: per tile-step cycles: MMA warp wait_a={v[0]:.0f} wait_full={v[1]:.0f} issue={v[2]:.0f} total={v[3]:.0f} | "
          f"epilogue(warp0) wait_d={v[4]:.0f} total={v[6]:.0f}")
