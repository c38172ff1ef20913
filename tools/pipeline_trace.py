"""Timeline of the pipelined plan loop (bulk on the main stream, refine/exchange on a side stream):
for each plan prints when its bulk half started/ended, when its result became ready, and the host
time spent issuing it.  Run alone (N=1) or under torch.distributed.run (sharded)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import FusedPlanner, synthetic  # noqa: E402


def main():
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    consume = int(os.environ.get("CONSUME", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sharded = None
    G, K, T, topk = 1, 65536, 8, 16
    gen, fwd = synthetic.make_models(nd=2, E=128, H=1024)
    pl = FusedPlanner(gen, fwd, horizon=T, device=dev, precision="f16+refine")
    if world > 1:
        import torch.distributed as dist
        from ndivplanning_b200.sharded import ShardedPlanner
        dist.init_process_group("nccl", device_id=dev)
        sharded = ShardedPlanner(pl)
    side = torch.cuda.Stream(device=dev, priority=-1)
    e0, goal, _ = synthetic.make_plan_inputs(G, 1, 1, 2, 128, seed=1)
    e0, goal = e0.to(dev), goal.to(dev)
    pool = [torch.rand(T, G, K, 2, device=dev) for _ in range(8)]

    ring = sharded.ring(G, K * world, topk) if sharded else pl.ring(G, K, topk)
    use_ring = int(os.environ.get("RING", "1"))

    def plan(nz):
        if use_ring:
            return ring.plan(e0, goal, nz)
        if sharded:
            return sharded.plan(e0, goal, nz, K * world, topk=topk, refine_stream=side)
        return pl.plan(e0, goal, nz, topk=topk, refine_stream=side)

    for i in range(3):
        plan(pool[i]).wait()
    torch.cuda.synchronize()
    n = 10
    ev = lambda: torch.cuda.Event(enable_timing=True)
    t0 = ev(); t0.record()
    rows, plans, host0 = [], [], time.perf_counter()
    for i in range(n):
        h0 = time.perf_counter()
        a = ev(); a.record()
        p = plan(pool[i % 8])
        # bulk end on main is not directly visible (the C call enqueues bulk then returns); record after the call
        b = ev(); b.record()
        r = ev(); r.record(ring.streams[(ring.count - 1) % len(ring.streams)] if use_ring else side)
        h1 = time.perf_counter()
        rows.append((a, b, r, (h0 - host0) * 1e3, (h1 - h0) * 1e3))
        plans.append(p)
        if consume and len(plans) > 2:
            plans.pop(0).wait()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"world={world} consume={consume}: plan | host issue at (ms) | host cost | main: start -> after-bulk | side ready")
        for i, (a, b, r, hs, hc) in enumerate(rows):
            print(f"  {i:2d} | {hs:7.3f} | {hc:6.3f} | {t0.elapsed_time(a):7.3f} -> {t0.elapsed_time(b):7.3f} | {t0.elapsed_time(r):7.3f}")
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
