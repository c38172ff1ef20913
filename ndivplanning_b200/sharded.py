"""Multi-GPU host logic (one process per GPU, torch.distributed for the plumbing).

MPC, one goal  : candidates are independent until the argmin (mpc_eval.py:125,159-165) -> rank r
                 scores candidates [r*K/P,(r+1)*K/P); ONE all-gather of a (4+topk)-word record per
                 goal (packed u64 argmin key okey(cost)<<32 | global index, the local winner's
                 action, verified flag, packed top-k keys) and a merge kernel select the plan with
                 the reference's lowest-index tie break.  No other data-path collective.
                 (allreduce_min_packed / merge_topk are the same reduction written with stock
                 collectives; the gloo tests use them to pin the key/tie-break semantics.)
MPC, many goals: shard goals, no collective.
ndiv           : rows are sharded in 128-aligned blocks; the [sx|sz] row sums are all-gathered between pass 1 and
                 pass 2 (the backward needs s_j for every j), then ONE all-gather of [grad rows | loss partial]
                 (ShardedNdiv; libndivplan's own communicator, caller's stream).
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

    @classmethod
    def solo(cls, device):
        """world-size-1 communicator without torch.distributed (single-GPU runs of the sharded entry points)"""
        import ctypes as C
        self = cls.__new__(cls)
        self.rank, self.world, self.device = 0, 1, torch.device(device)
        idb = (C.c_ubyte * 128)()
        check(lib().ndp_comm_unique_id(idb))
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(lib().ndp_comm_create(C.byref(self._h), idb, 0, 1))
        return self

    def allgather(self, send: torch.Tensor) -> torch.Tensor:
        """rank-major all-gather on the current stream: [..] -> [world, ..]"""
        send = send.contiguous()
        out = torch.empty((self.world,) + tuple(send.shape), dtype=send.dtype, device=send.device)
        with torch.cuda.device(self.device):
            check(lib().ndp_comm_allgather_bytes(self._h, ptr(send), ptr(out), send.numel() * send.element_size(),
                                                 stream_ptr(self.device)))
        return out

    def peer_status(self) -> bool:
        """True when the exchanges run over the mapped peer window (NVLink loads/stores + flags) instead of NCCL;
        raises if an in-kernel wait for a peer expired since the last call.  Synchronises the current stream."""
        import ctypes as C
        m = C.c_int(0)
        with torch.cuda.device(self.device):
            check(lib().ndp_comm_peer_status(self._h, C.byref(m), stream_ptr(self.device)))
        return bool(m.value)

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


def exchange_plan(p: Plan, topk: int, comm: "Comm"):
    """local plan -> global plan on the current stream through ndp_shard_exchange (ONE library call: pack, exchange
    over the peer window — or ncclAllGather — and merge).  Same results as pack_records -> exchange_records ->
    merge_records."""
    G = p.packed_min.shape[0]
    dev, W, P = p.packed_min.device, 4 + topk, comm.world
    i64, i32, f32 = torch.int64, torch.int32, torch.float32
    rec = torch.empty(G, W, dtype=i64, device=dev)
    rec_all = torch.empty(P, G, W, dtype=i64, device=dev)
    packed, best_idx = torch.empty(G, dtype=i64, device=dev), torch.empty(G, dtype=i32, device=dev)
    best_action, verified = torch.empty(G, 4, dtype=f32, device=dev), torch.empty(G, dtype=i32, device=dev)
    tk_idx = torch.empty(G, topk, dtype=i32, device=dev) if topk else None
    tk_cost = torch.empty(G, topk, dtype=f32, device=dev) if topk else None
    with torch.cuda.device(dev):
        check(lib().ndp_shard_exchange(comm._h, ptr(p.packed_min), ptr(p.best_action), ptr(p.verified), ptr(p.topk_idx),
                                       ptr(p.topk_cost), G, topk, ptr(rec), ptr(rec_all), ptr(packed), ptr(best_idx),
                                       ptr(best_action), ptr(tk_idx), ptr(tk_cost), ptr(verified), stream_ptr(dev)))
    return packed, best_idx, best_action, tk_idx, tk_cost, verified


class ShardedPlanner:
    """Candidate-sharded planning: every rank holds the weights, e0 and goal; noise is per shard.
    Per plan: local K1/K3 (+refine), then ndp_shard_exchange: the pack kernel stores the (4+topk)*8-byte record per goal
    into every rank's peer window over NVLink, the merge kernel waits for every rank's flag (ncclAllGather between
    ndp_shard_pack and ndp_shard_merge when the ranks cannot map each other).  With `refine_stream` the refine half, the exchange and the merge all run
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
            packed, best_idx, best_action, tk_idx, tk_cost, verified = exchange_plan(p, topk, self.comm)
            ready = None
            if side is not None:
                ready = torch.cuda.Event()
                ready.record(side)
        if side is not None:                       # results are produced on `side`, consumed on the caller's stream
            for t_ in (best_idx, best_action, packed, tk_idx, tk_cost, verified):
                if t_ is not None:
                    t_.record_stream(torch.cuda.current_stream(dev))
        return Plan(best_idx, best_action, packed, p.cost, p.a0, tk_idx, tk_cost, p.state_T, verified, ready)


def ndiv_row_block(k: int, rank: int, world: int):
    """row block of rank `rank`: equal blocks rounded up to 128 rows (the Gram kernels' tile; keeps the tensor-core
    path), the last ranks may own fewer rows or none.  Returns (row0, rows, block)."""
    block = (-(-k // world) + 127) // 128 * 128
    row0 = min(k, rank * block)
    return row0, min(k, row0 + block) - row0, block


class ShardedNdiv:
    """Row-sharded ndiv loss forward + backward (reference diversity.py:12-19,36-41): x, z are
    replicated, rank r owns a 128-aligned row block.  Two collectives per call, both on libndivplan's own NCCL
    communicator and on the CALLER's stream (CUDA-graph capturable, no torch.distributed in the data path):

        pass 1 (ndp_ndiv_rowsums, local rows)      -> all-gather of the rank's [sx | sz] block  (8*N*block bytes)
        pass 2 (ndp_ndiv_loss_grad, local rows)    -> all-gather of [grad_x rows | loss partial] (4*N*block*Dx + 4)

    The loss is the sum of the gathered partials in rank order, so every rank holds the same bits (no all-reduce
    rounding-order ambiguity) and — row blocks being 128-aligned — the same bits as the single-GPU call.  Buffers are
    allocated once per (N, k, Dx, Dz).

    One group on the Gram path (N == 1, k >= 1536, Dx, Dz <= 32) takes the FUSED entry instead: one C call per rank
    (ndp_ndiv_sharded_forward_backward: pre-pack once, pass kernels writing straight into the exchange slots, in-place
    all-gathers, two unpack kernels; column split sized to the rank's rows so 1/P of the rows still fills the GPU).
    Its results are identical on every rank and equal the single-GPU call to fp32 round-off of the split order; the
    returned tensors are owned by this object and overwritten by the next call."""

    def __init__(self, comm: Comm, N: int, k: int, Dx: int, Dz: int, fused=None):
        import ctypes as C
        self.comm, self.N, self.k, self.Dx, self.Dz = comm, N, k, Dx, Dz
        dev = comm.device
        self.row0, self.rows, self.block = ndiv_row_block(k, comm.rank, comm.world)
        W, B, f32 = comm.world, self.block, torch.float32
        self.sx = torch.zeros(N, W * B, dtype=f32, device=dev)          # padded to W*B rows; [:, :k] is the result
        self.sz = torch.zeros(N, W * B, dtype=f32, device=dev)
        self.s_send = torch.zeros(2, N, B, dtype=f32, device=dev)
        self.s_recv = torch.empty(W, 2, N, B, dtype=f32, device=dev)
        self.g_send = torch.zeros(N * B * Dx + 1, dtype=f32, device=dev)
        self.g_recv = torch.empty(W, N * B * Dx + 1, dtype=f32, device=dev)
        self.grad = torch.zeros(N, k, Dx, dtype=f32, device=dev)        # local rows written by pass 2
        self.loss_part = torch.zeros(1, dtype=f32, device=dev)
        n = C.c_size_t()
        check(lib().ndp_ndiv_workspace_bytes(N, k, Dx, Dz, C.byref(n)))
        self.ws = torch.empty(max(n.value, 256), dtype=torch.uint8, device=dev)
        # one group on the tensor-core (Gram) path: ONE library call per rank does pre-pack, both passes and both
        # exchanges with the kernels writing straight into the exchange slots (ndp_ndiv_sharded_forward_backward)
        self.fused = fused if fused is not None else (N == 1 and k >= 1536 and Dx <= 32 and Dz <= 32 and Dx % 4 == 0)
        if self.fused:
            check(lib().ndp_ndiv_sharded_workspace_bytes(k, Dx, Dz, W, C.byref(n)))
            self.fws = torch.empty(n.value + 256, dtype=torch.uint8, device=dev)
            self.fws = self.fws[(-self.fws.data_ptr()) % 256:][:n.value]
            self.f_sx = torch.empty(k, dtype=f32, device=dev)
            self.f_sz = torch.empty(k, dtype=f32, device=dev)
            self.f_loss = torch.empty(1, dtype=f32, device=dev)
            self.f_grad = torch.empty(1, k, Dx, dtype=f32, device=dev)

    def __call__(self, x: torch.Tensor, z: torch.Tensor, want_grad: bool = True):
        """x [N,k,Dx], z [N,k,Dz] on this rank's device -> (loss[()], grad_x [N,k,Dx] | None), complete on every rank"""
        N, k, Dx, Dz, B, W = self.N, self.k, self.Dx, self.Dz, self.block, self.comm.world
        dev = self.comm.device
        L, st = lib(), stream_ptr(dev)
        r0, rows = self.row0, self.rows
        if self.fused:
            with torch.cuda.device(dev):
                check(L.ndp_ndiv_sharded_forward_backward(self.comm._h, ptr(x), ptr(z), k, Dx, Dz, _div.MARGIN,
                                                          ptr(self.f_sx), ptr(self.f_sz), ptr(self.f_loss),
                                                          ptr(self.f_grad) if want_grad else None, ptr(self.fws),
                                                          self.fws.numel(), st))
            return self.f_loss.reshape(()), (self.f_grad if want_grad else None)
        with torch.cuda.device(dev):
            if rows > 0:
                sx_l = torch.empty(N, k, dtype=torch.float32, device=dev)
                sz_l = torch.empty(N, k, dtype=torch.float32, device=dev)
                check(L.ndp_ndiv_rowsums(ptr(x), ptr(z), N, k, Dx, Dz, r0, rows, ptr(sx_l), ptr(sz_l), ptr(self.ws),
                                         self.ws.numel(), st))
                self.s_send[0, :, :rows].copy_(sx_l[:, r0:r0 + rows])
                self.s_send[1, :, :rows].copy_(sz_l[:, r0:r0 + rows])
            check(L.ndp_comm_allgather_f32(self.comm._h, ptr(self.s_send), ptr(self.s_recv), self.s_send.numel(), st))
            sx = self.s_recv[:, 0].permute(1, 0, 2).reshape(N, W * B)[:, :k].contiguous()
            sz = self.s_recv[:, 1].permute(1, 0, 2).reshape(N, W * B)[:, :k].contiguous()
            if rows > 0:
                check(L.ndp_ndiv_loss_grad(ptr(x), ptr(z), ptr(sx), ptr(sz), N, k, Dx, Dz, _div.MARGIN, r0, rows,
                                           ptr(self.loss_part), ptr(self.grad) if want_grad else None, ptr(self.ws),
                                           self.ws.numel(), st))
                if want_grad:
                    self.g_send[:-1].view(N, B, Dx)[:, :rows].copy_(self.grad[:, r0:r0 + rows])
                self.g_send[-1:].copy_(self.loss_part)
            else:
                self.g_send.zero_()
            cnt = self.g_send.numel() if want_grad else 1
            send = self.g_send if want_grad else self.g_send[-1:]
            recv = self.g_recv if want_grad else self.g_recv.view(-1)[:W]
            check(L.ndp_comm_allgather_f32(self.comm._h, ptr(send), ptr(recv), cnt, st))
            if want_grad:
                parts = self.g_recv[:, -1]
                grad = self.g_recv[:, :-1].view(W, N, B, Dx).permute(1, 0, 2, 3).reshape(N, W * B, Dx)[:, :k].contiguous()
            else:
                parts, grad = recv, None
            loss = parts[0].clone()
            for r in range(1, W):                       # fixed (rank) order: identical bits on every rank
                loss = loss + parts[r]
        return loss.reshape(()), grad


def sharded_ndiv_loss(x: torch.Tensor, z: torch.Tensor, comm: Comm, want_grad: bool = True):
    """one-shot convenience around ShardedNdiv (allocates its buffers per call)"""
    N, k, Dx = x.shape
    return ShardedNdiv(comm, N, k, Dx, z.shape[2])(x, z, want_grad)
