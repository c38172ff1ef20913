"""Dev: cycle counters of the two-tile rollout kernel (needs the -DNDP_TC_PROF build: tools/build_prof.sh, run with
NDP_LIB_PATH=ndivplanning_b200/libndivplan_prof.so)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ndivplanning_b200 import FusedPlanner, synthetic, _lib
dev = torch.device("cuda", 0)
out = (C.c_ulonglong * 96)()
L = C.CDLL(_lib.LIB_PATH)
for (G, K, T, H) in ((1, 65536, 8, 1024), (1, 65536, 8, 256), (1, 37888, 8, 1024), (1, 37888, 1, 1024)):
    gen, fwd = synthetic.make_models(nd=2, E=128, H=H)
    e0, goal, noise = [x.to(dev) for x in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=5)]
    pl = FusedPlanner(gen, fwd, horizon=T, precision="f16", device=dev)
    pl.rollout_score(e0, goal, noise); torch.cuda.synchronize()
    L.ndp_debug_tc_prof(out, 1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); pl.rollout_score(e0, goal, noise); b.record(); torch.cuda.synchronize()
    L.ndp_debug_tc_prof(out, 1)
    n = ((K + 255) // 256) * T          # pair-steps
    v = [out[88 + i] / n for i in range(8)]
    raw = [out[88 + i] for i in range(8)]
    ctas = max(raw[3], 1)
    print(f"  CTAs {raw[3]}: life {raw[7] / ctas:.0f} cycles = {raw[6] / ctas / 1e3:.1f} us per CTA -> effective SM clock "
          f"{raw[7] / max(raw[6], 1):.3f} GHz; loop cycles per CTA {raw[2] / ctas * 1.0:.0f} (issuer) -> outside the step "
          f"loop {(raw[7] - raw[2]) / ctas:.0f} cycles per CTA", flush=True)
    print(f"K={K} H={H}: {a.elapsed_time(b):.4f} ms | per pair-step cycles: issuer(tile0) wait_a={v[0]:.0f} wait_full={v[1]:.0f} "
          f"total={v[2]:.0f} | epilogue(warp0) wait_d={v[4]:.0f} total={v[5]:.0f}", flush=True)
