"""Multi-GPU host logic (one process per GPU, torch.distributed for the plumbing).

MPC, one goal  : candidates are independent until the argmin (mpc_eval.py:125,159-165) -> rank r
                 scores candidates [r*K/P,(r+1)*K/P); ONE all-gather of a (4+topk)-word record per
                 goal (packed u64 argmin key okey(cost)<<32 | global index, the local winner's
                 action, verified flag, packed top-k keys) and a merge kernel select the plan with
                 the reference's lowest-index tie break.  No other data-path collective.
                 (allreduce_min_packed / merge_topk are the same reduction written with stock
                 collectives; the gloo tests use them to pin the key/tie-break semantics.)
MPC, many goals: shard goals, no collective.
ndiv           : rows are sharded; row sums are all-gathered between pass 1 and pass 2
                 (the backward needs s_j for every j), loss/grad partials are summed.
The collectives are latency-bound (8 B .. 0.5 MB).
"""
import torch
import torch.distributed as dist

from . import diversity as _div
from ._lib import NdpError, check, lib, ptr, stream_ptr
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


def pack_records(p: Plan, topk: int) -> torch.Tensor:
    """local plan -> [G, 4+topk] int64 records (ndp_shard_pack, include/ndivplan.h)"""
    G = p.packed_min.shape[0]
    dev = p.packed_min.device
    rec = torch.empty(G, 4 + topk, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        check(lib().ndp_shard_pack(ptr(p.packed_min), ptr(p.best_action), ptr(p.verified), ptr(p.topk_idx),
                                   ptr(p.topk_cost), G, topk, ptr(rec), stream_ptr(dev)))
    return rec


class Comm:
    """libndivplan's own NCCL communicator (csrc/comm.cu): collectives are enqueued on the CALLER's
    stream — here the high-priority refine stream — instead of torch's internal NCCL stream.  The
    128-byte ncclUniqueId is created on rank 0 and broadcast through torch.distributed (plumbing)."""

    def __init__(self, device, group=None):
        import ctypes as C
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = torch.device(device)
        idb = (C.c_ubyte * 128)()
        if self.rank == 0:
            check(lib().ndp_comm_unique_id(idb))
        t = torch.tensor(list(idb), dtype=torch.uint8, device=self.device)
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        idb = (C.c_ubyte * 128)(*t.cpu().tolist())
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().ndp_comm_create(C.byref(self._h), idb, self.rank, self.world))

    def allgather(self, send: torch.Tensor) -> torch.Tensor:
        """rank-major all-gather on the current stream: [..] -> [world, ..]"""
        send = send.contiguous()
        out = torch.empty((self.world,) + tuple(send.shape), dtype=send.dtype, device=send.device)
        with torch.cuda.device(self.device):
            check(lib().ndp_comm_allgather_bytes(self._h, ptr(send), ptr(out), send.numel() * send.element_size(),
                                                 stream_ptr(self.device)))
        return out

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().ndp_comm_destroy(self._h)
            self._h = None


def exchange_records(rec: torch.Tensor, group=None, comm: "Comm" = None) -> torch.Tensor:
    """the ONE collective of a sharded plan: all-gather of the per-rank records -> [P,G,W]"""
    if comm is not None:
        return comm.allgather(rec)
    world = dist.get_world_size(group)
    out = torch.empty((world,) + tuple(rec.shape), dtype=rec.dtype, device=rec.device)
    if rec.is_cuda:
        dist.all_gather_into_tensor(out, rec.contiguous(), group=group)
    else:                                            # gloo (host-logic tests) has no flat all-gather
        dist.all_gather(list(out.unbind(0)), rec.contiguous(), group=group)
    return out


def merge_records(rec_all: torch.Tensor, topk: int):
    """[P,G,4+topk] records -> (packed[G], best_idx[G], best_action[G,4], topk_idx, topk_cost, verified[G])
    (ndp_shard_merge)"""
    P, G, _ = rec_all.shape
    dev = rec_all.device
    packed = torch.empty(G, dtype=torch.int64, device=dev)
    best_idx = torch.empty(G, dtype=torch.int32, device=dev)
    best_action = torch.empty(G, 4, dtype=torch.float32, device=dev)
    verified = torch.empty(G, dtype=torch.int32, device=dev)
    tk_idx = torch.empty(G, topk, dtype=torch.int32, device=dev) if topk else None
    tk_cost = torch.empty(G, topk, dtype=torch.float32, device=dev) if topk else None
    with torch.cuda.device(dev):
        check(lib().ndp_shard_merge(ptr(rec_all), P, G, topk, ptr(packed), ptr(best_idx), ptr(best_action),
                                    ptr(tk_idx), ptr(tk_cost), ptr(verified), stream_ptr(dev)))
    return packed, best_idx, best_action, tk_idx, tk_cost, verified


class ShardedPlanner:
    """Candidate-sharded planning: every rank holds the weights, e0 and goal; noise is per shard.
    Per plan: local K1/K3 (+refine), ndp_shard_pack, ONE all-gather of (4+topk)*8 bytes per goal and
    rank, ndp_shard_merge.  With `refine_stream` the refine half, the exchange and the merge all run
    on that stream under the bulk rollout of the next plan (Plan.ready / Plan.wait())."""

    def __init__(self, planner: FusedPlanner, group=None):
        self.planner, self.group = planner, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.comm = Comm(planner.device, group)

    def ring(self, G: int, K_total: int, topk: int = 0, depth: int = 4, n_streams: int = 2, fallback: bool = False):
        """PlanRing over this rank's shard of K_total candidates: allocation-free steady-state loop with the
        refine tail, the record exchange (on libndivplan's own communicator) and the merge on side streams"""
        start, count = shard_range(K_total, self.rank, self.world)
        return self.planner.ring(G, count, topk, depth, n_streams, idx_base=start, comm=self.comm, fallback=fallback)

    def plan(self, e0, goal, noise_local, K_total: int, topk: int = 0, refine_stream=None) -> Plan:
        start, count = shard_range(K_total, self.rank, self.world)
        assert noise_local.shape[2] == count, (noise_local.shape, count)
        if self.world * topk > 1024:
            raise NdpError("sharded top-k needs world*topk <= 1024")
        p = self.planner.plan(e0, goal, noise_local, topk=topk, idx_base=start, refine_stream=refine_stream)
        dev = p.cost.device
        side = refine_stream if p.ready is not None else None
        with torch.cuda.stream(side if side is not None else torch.cuda.current_stream(dev)):
            rec_all = exchange_records(pack_records(p, topk), self.group, self.comm)
            packed, best_idx, best_action, tk_idx, tk_cost, verified = merge_records(rec_all, topk)
            ready = None
            if side is not None:
                ready = torch.cuda.Event()
                ready.record(side)
        if side is not None:                       # results are produced on `side`, consumed on the caller's stream
            for t_ in (best_idx, best_action, packed, tk_idx, tk_cost, verified):
                if t_ is not None:
                    t_.record_stream(torch.cuda.current_stream(dev))
        return Plan(best_idx, best_action, packed, p.cost, p.a0, tk_idx, tk_cost, p.state_T, verified, ready)


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
