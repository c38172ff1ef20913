"""Multi-GPU host logic (one process per GPU, torch.distributed for the plumbing).

MPC, one goal  : candidates are independent until the argmin (mpc_eval.py:125,159-165) -> rank r
                 scores candidates [r*K/P,(r+1)*K/P); ONE min-allreduce of the packed u64
                 (okey(cost)<<32 | global index) selects the plan with the reference's
                 lowest-index tie break.  No other data-path collective.
MPC, many goals: shard goals, no collective.
ndiv           : rows are sharded; row sums are all-gathered between pass 1 and pass 2
                 (the backward needs s_j for every j), loss/grad partials are summed.
The collectives are latency-bound (8 B .. 0.5 MB).
"""
import torch
import torch.distributed as dist

from . import diversity as _div
from .planner import FusedPlanner, Plan, select_plan

_SIGN = -0x8000000000000000  # int64 min: xor flips the top bit


def shard_range(total: int, rank: int, world: int):
    """contiguous block partition; the first (total % world) ranks get one extra unit"""
    base, rem = divmod(total, world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def allreduce_min_packed(packed: torch.Tensor, group=None) -> torch.Tensor:
    """unsigned-64 min over ranks of int64-viewed packed keys.  torch.distributed has no
    uint64 reductions; flipping the top bit maps unsigned order onto signed order exactly."""
    t = packed ^ _SIGN
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return t ^ _SIGN


def merge_topk(tk_cost: torch.Tensor, tk_idx: torch.Tensor, topk: int, group=None, select=None):
    """all-gather every rank's ascending top-k list and keep the global top-k.
    tk_cost/tk_idx [G,topk] -> ([G,topk] cost, [G,topk] global idx).  Rank-major gather order
    equals global index order inside cost ties (ranks own increasing index ranges)."""
    world = dist.get_world_size(group)
    costs = [torch.empty_like(tk_cost) for _ in range(world)]
    idxs = [torch.empty_like(tk_idx) for _ in range(world)]
    dist.all_gather(costs, tk_cost.contiguous(), group=group)
    dist.all_gather(idxs, tk_idx.contiguous(), group=group)
    allc, alli = torch.cat(costs, dim=1), torch.cat(idxs, dim=1)          # [G, world*topk]
    allc = torch.where(alli < 0, torch.full_like(allc, float("nan")), allc)   # padded entries sort last
    if select is None:
        pos = select_plan(allc, topk)[1].long()                          # K3 on the gathered keys
    else:
        pos = select(allc, topk)
    pos = pos.clamp_min(0)
    return torch.gather(allc, 1, pos), torch.gather(alli, 1, pos)


class ShardedPlanner:
    """Candidate-sharded planning: every rank holds the weights, e0 and goal; noise is per shard."""

    def __init__(self, planner: FusedPlanner, group=None):
        self.planner, self.group = planner, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def plan(self, e0, goal, noise_local, K_total: int, topk: int = 0) -> Plan:
        start, count = shard_range(K_total, self.rank, self.world)
        assert noise_local.shape[2] == count, (noise_local.shape, count)
        p = self.planner.plan(e0, goal, noise_local, topk=topk, idx_base=start)
        packed = allreduce_min_packed(p.packed_min, self.group)            # the ONE collective per plan
        mine = (packed == p.packed_min) & (packed != -0x100000000)         # owner of the winning candidate
        # winner's action: owner contributes it, everybody else zeros (16 B per goal)
        best_action = torch.where(mine[:, None], p.best_action, torch.zeros_like(p.best_action))
        # all-invalid plan decodes to global index 0 (mpc_eval.py:122,159): rank 0 owns it
        if self.rank == 0:
            inval = packed == -0x100000000
            best_action = torch.where(inval[:, None], p.a0[:, 0, :], best_action)
        dist.all_reduce(best_action, group=self.group)
        best_idx = (packed & 0xFFFFFFFF).to(torch.int32)
        tk_cost = tk_idx = None
        if topk:
            tk_cost, tk_idx = merge_topk(p.topk_cost, p.topk_idx, topk, self.group)
        return Plan(best_idx, best_action, packed, p.cost, p.a0, tk_idx, tk_cost, p.state_T)


def sharded_ndiv_loss(x: torch.Tensor, z: torch.Tensor, group=None, want_grad: bool = True):
    """Row-sharded ndiv forward+backward: x [N,k,Dx], z [N,k,Dz] replicated on every rank.
    Returns (loss[()] summed over all rows, grad_x [N,k,Dx] complete on every rank)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    N, k, _ = x.shape
    row0, rows = shard_range(k, rank, world)
    sx = torch.zeros(N, k, dtype=torch.float32, device=x.device)
    sz = torch.zeros_like(sx)
    _div.ndiv_rowsums(x, z, row0, rows, sx, sz)
    s = torch.stack([sx, sz])                        # rows outside the block are 0 -> sum == gather
    dist.all_reduce(s, group=group)                  # 2*N*k*4 bytes (32 KB at k=4096)
    loss, grad = _div.ndiv_loss_grad(x, z, s[0].contiguous(), s[1].contiguous(), want_grad, row0, rows)
    dist.all_reduce(loss, group=group)
    if want_grad:
        dist.all_reduce(grad, group=group)           # row blocks are disjoint: sum == gather
    return loss.reshape(()), grad
