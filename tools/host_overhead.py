import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ndivplanning_b200 import FusedPlanner, synthetic
dev = torch.device("cuda", 0)
G, K, T, H = 1, 65536, 8, 1024
gen, fwd = synthetic.make_models(nd=2, E=128, H=H)
e0, goal, noise = [x.to(dev) for x in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=5)]
side = torch.cuda.Stream(device=dev, priority=-1)
for mode, rs in [("f16+refine", side), ("f16+refine", None), ("f16", None)]:
    pl = FusedPlanner(gen, fwd, horizon=T, precision=mode, device=dev)
    for _ in range(3): pl.plan(e0, goal, noise, topk=16, refine_stream=rs).wait()
    torch.cuda.synchronize()
    n = 30
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(n):
        last = pl.plan(e0, goal, noise, topk=16, refine_stream=rs)
    t1 = time.perf_counter()
    last.wait(); b.record(); torch.cuda.synchronize()
    print(f"{mode} side={rs is not None}: host {1e3*(t1-t0)/n:.3f} ms/plan, device {a.elapsed_time(b)/n:.3f} ms/plan")

# ---- split host timing of the two ABI halves
import ctypes as C
from ndivplanning_b200._lib import lib, ptr, stream_ptr
pl = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine", device=dev)
cost = torch.empty(G, K, device=dev); a0 = torch.empty(G, K, 4, device=dev)
packed = torch.empty(G, dtype=torch.int64, device=dev); ver = torch.empty(G, dtype=torch.int32, device=dev)
tki = torch.empty(G, 16, dtype=torch.int32, device=dev); tkc = torch.empty(G, 16, device=dev)
n = C.c_size_t(); lib().ndp_plan_workspace_bytes(pl._h, G, K, 16, C.byref(n))
ws = torch.empty(n.value, dtype=torch.uint8, device=dev)
args = (pl._h, ptr(e0), ptr(goal), ptr(noise), G, K, 16, 0, 0.0, ptr(cost), ptr(a0), ptr(packed), ptr(tki), ptr(tkc), ptr(ver), ptr(ws), ws.numel(), stream_ptr(dev))
for name in ("ndp_plan_select_bulk", "ndp_plan_select_refine"):
    fn = getattr(lib(), name)
    for _ in range(3): fn(*args)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20): fn(*args)
    t1 = time.perf_counter(); torch.cuda.synchronize()
    print(name, f"host {1e3*(t1-t0)/20:.3f} ms/call")
