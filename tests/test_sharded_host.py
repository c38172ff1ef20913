"""CPU, world_size=2 over gloo: the host-side logic of the sharded path (index ranges, the
unsigned-min allreduce of packed keys, top-k merge).  Kernels are not involved."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ndiv_oracle as O


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _to_i64(u):
    return u - (1 << 64) if u >= (1 << 63) else u


def _worker(rank, world, port, K, topk):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ndivplanning_b200.sharded import allreduce_min_packed, merge_topk, shard_range
        torch.manual_seed(7)
        cost = torch.rand(3, K)
        cost[1, :] = float("nan")                 # goal 1: nothing valid -> invalid key -> index 0
        cost[2, 5] = cost[2, K - 3] = -1.0        # goal 2: tie across ranks -> lowest index wins
        start, count = shard_range(K, rank, world)
        assert sum(shard_range(K, r, world)[1] for r in range(world)) == K
        local = cost[:, start:start + count]
        packed = []
        for g in range(3):
            i = O.select_first_min(local[g])
            valid = bool((local[g] < 1e10).any())
            packed.append(_to_i64(O.pack_key(float(local[g, i]), start + i) if valid else 0xFFFFFFFF00000000))
        got = allreduce_min_packed(torch.tensor(packed, dtype=torch.int64))
        for g in range(3):
            want = O.select_first_min(cost[g])
            valid = bool((cost[g] < 1e10).any())
            assert int(got[g]) & 0xFFFFFFFF == (want if valid else 0), (g, int(got[g]) & 0xFFFFFFFF, want)
        assert int(got[2]) & 0xFFFFFFFF == 5
        # top-k merge (selection function = oracle stable sort on CPU)
        order = torch.sort(local, dim=1, stable=True).indices[:, :topk]
        tk_cost = torch.gather(local, 1, order)
        tk_idx = (order + start).to(torch.int32)
        mc, mi = merge_topk(tk_cost, tk_idx, topk,
                            select=lambda c, k: torch.sort(c, dim=1, stable=True).indices[:, :k])
        for g in (0, 2):
            assert mi[g].tolist() == O.select_topk(cost[g], topk).tolist()
        # the product's single collective: all-gather of per-rank records, rank-major
        from ndivplanning_b200.sharded import exchange_records
        rec = torch.full((3, 4 + topk), rank, dtype=torch.int64) + torch.arange(4 + topk)
        allrec = exchange_records(rec)
        assert allrec.shape == (world, 3, 4 + topk)
        for r in range(world):
            assert torch.equal(allrec[r], torch.full((3, 4 + topk), r, dtype=torch.int64) + torch.arange(4 + topk))
    finally:
        dist.destroy_process_group()


def test_sharded_host_logic_gloo_world2():
    mp.spawn(_worker, args=(2, _free_port(), 1001, 8), nprocs=2, join=True)
