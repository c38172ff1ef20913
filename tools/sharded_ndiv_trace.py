"""Per-kernel device times of the row-sharded ndiv step on N ranks (torch.profiler / CUPTI activity records: no kernel
replay, so the cross-rank flag waits behave as in a normal run).  Launch under torch.distributed.run like
tools/sharded_check.py; rank 0 prints the table."""
import os
import sys

import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import synthetic  # noqa: E402
from ndivplanning_b200.sharded import Comm, ShardedNdiv  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = Comm(dev)
    x, z = synthetic.make_ndiv_inputs(1, 4096, 32, 16)
    x, z = x.to(dev), z.to(dev)
    op = ShardedNdiv(comm, 1, 4096, 32, 16)
    align = torch.zeros(1, device=dev)
    for _ in range(5):
        op(x, z)
    torch.cuda.synchronize()
    dist.barrier()
    steps = 20
    evs = []
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(steps):
            torch.cuda._sleep(1_000_000)
            dist.all_reduce(align)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); op(x, z); b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    if rank == 0:
        print(f"world {world}: step {ms * 1e3:.1f} us (events), peer window {comm.peer_status()}")
        rows = [(e.key, e.device_time_total / max(e.count, 1), e.count) for e in prof.key_averages()
                if e.device_time_total > 0]
        for k_, t, c in sorted(rows, key=lambda r: -r[1]):
            print(f"  {t:8.1f} us x{c:4d}  {k_[:110]}")
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
