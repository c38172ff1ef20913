"""FusedPlanner — the MPC inner loop of the reference (mpc_eval.py:124-165) as three kernels:
K1 fused generator + T-step rollout + terminal cost, K3 segmented argmin/top-k, gather.

    planner = FusedPlanner(generator, fwd_model, horizon=8)
    out = planner.plan(e0[G,E], goal[G,E], noise[T,G,K,nd], topk=16)
    out.best_idx[G]  out.best_action[G,4]  out.topk_idx[G,topk]  out.cost[G,K]

`generator` is anything with the state_dict of reference models.gan.Decoder (fc1..fc5),
`fwd_model` anything with LatentForward's (fc1, fc2).  CUDA-only; no CPU fallback.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import PlannerDesc, check, f32c, lib, ptr, require_cuda, stream_ptr

INVALID_PACKED = 0xFFFFFFFF00000000


@dataclass
class Plan:
    best_idx: torch.Tensor        # [G] int32 (global candidate index)
    best_action: torch.Tensor     # [G,4]
    packed_min: torch.Tensor      # [G] int64 view of the u64 (okey(cost)<<32 | idx)
    cost: torch.Tensor            # [G,K]
    a0: torch.Tensor              # [G,K,4]
    topk_idx: Optional[torch.Tensor] = None    # [G,topk] int32
    topk_cost: Optional[torch.Tensor] = None   # [G,topk]
    state_T: Optional[torch.Tensor] = None     # [G,K,E]
    verified: Optional[torch.Tensor] = None    # [G] int32: 1 = selection proven equal to the exact-cost selection
    ready: Optional[torch.cuda.Event] = None   # set when the plan was finished on a side stream (refine_stream=)
    _info: Optional[tuple] = None              # (planner handle, G, K, topk, workspace) for verify_info()

    def verify_info(self):
        """[G,2] float32 {margin, eta} of the refine proof in sqrt-cost units (ndp_plan_verify_info): margin > 0 <=>
        verified == 1; eta = the error bound assumed for the single-pass estimate.  None for non-refined plans."""
        if self._info is None:
            return None
        h, G, K, topk, ws = self._info
        dev = self.cost.device
        self.wait()
        out = torch.empty(G, 2, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib().ndp_plan_verify_info(h, G, K, topk, ptr(ws), ws.numel(), ptr(out), stream_ptr(dev)))
        return out

    def wait(self, stream=None):
        """make `stream` (default: current) wait until the plan's tensors are final"""
        if self.ready is not None:
            (stream or torch.cuda.current_stream(self.cost.device)).wait_event(self.ready)
        return self


def select_plan(cost: torch.Tensor, topk: int = 0, idx_base: int = 0):
    """K3 alone: cost [G,K] -> (packed_min[G] int64, topk_idx[G,topk] int32 | None, topk_cost | None).
    Replaces the python scan mpc_eval.py:159-164 (strict <, first index, init 1e10, NaN loses)."""
    require_cuda(cost)
    c = f32c(cost)
    G, K = c.shape
    dev = c.device
    packed = torch.empty(G, dtype=torch.int64, device=dev)
    tk_idx = torch.empty(G, topk, dtype=torch.int32, device=dev) if topk > 0 else None
    tk_cost = torch.empty(G, topk, dtype=torch.float32, device=dev) if topk > 0 else None
    n = C.c_size_t()
    check(lib().ndp_select_workspace_bytes(G, K, topk, C.byref(n)))
    ws = torch.empty(max(n.value, 8), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib().ndp_select(ptr(c), G, K, topk, idx_base, ptr(packed), ptr(tk_idx), ptr(tk_cost), ptr(ws),
                               ws.numel(), stream_ptr(dev)))
    return packed, tk_idx, tk_cost


def unpack_index(packed: torch.Tensor) -> torch.Tensor:
    """index half of the packed key (invalid key decodes to 0, like best_act_ind = 0)."""
    return (packed & 0xFFFFFFFF).to(torch.int64)


class FusedPlanner:
    def __init__(self, generator, fwd_model, *, horizon: int, precision: Optional[str] = None, device=None):
        """precision: None (default) = "f16+refine" — tcgen05 bulk estimate of all K, split-operand re-roll of the 128
        best, data-derived proof per goal, exact re-roll of the goals that cannot be proven (`plan(fallback=True)`) —
        whenever the shape is one the tensor-core kernels take (E=128, noise_dim<=15, H <= 2048; any hidden
        width — the library pads it to a multiple of 128 with zero units, exactly), else "fp32".  "fp32" (CUDA-core, every cost fp32), "f16x3", "f16x2", "f16" by request."""
        g, f = generator.state_dict(), fwd_model.state_dict()
        names = ["fc1", "fc2", "fc3", "fc4", "fc5"]
        Ws = [g[f"{n}.weight"] for n in names] + [f["fc1.weight"], f["fc2.weight"]]
        bs = [g[f"{n}.bias"] for n in names] + [f["fc1.bias"], f["fc2.bias"]]
        H, EA = f["fc1.weight"].shape
        E = f["fc2.weight"].shape[0]
        if EA != E + 4:
            raise _lib.NdpError(f"forward model fc1 expects E+4 inputs, got {EA} for E={E}")
        nd = g["fc1.weight"].shape[1] - 2 * E
        if nd < 0 or tuple(g["fc5.weight"].shape) != (4, 256):
            raise _lib.NdpError("generator does not have the models.gan.Decoder layout (gan.py:67-71)")
        if device is None:
            device = Ws[0].device if Ws[0].is_cuda else torch.device("cuda", torch.cuda.current_device()) \
                if torch.cuda.is_available() else None
        if device is None:
            raise _lib.NdpError("CUDA is not available; FusedPlanner has no CPU fallback")
        self.device = torch.device(device)
        self.E, self.H, self.nd, self.T = E, H, nd, int(horizon)
        if precision is None:
            precision = "f16+refine" if (E == 128 and 0 <= nd <= 15 and 1 <= H <= 2048) else "fp32"
        prec = {"fp32": _lib.PREC_FP32, "f16x3": _lib.PREC_F16X3, "f16x2": _lib.PREC_F16X2,
                "f16": _lib.PREC_F16, "f16+refine": _lib.PREC_F16_REFINE}[precision]
        self.precision = precision
        self._h = C.c_void_p()
        desc = PlannerDesc(E, nd, 4, H, self.T, prec)
        with torch.cuda.device(self.device):
            check(lib().ndp_planner_create(C.byref(self._h), C.byref(desc)))
            Wd = [f32c(w.to(self.device)) for w in Ws]
            bd = [f32c(b.to(self.device)) for b in bs]
            Wp = (C.c_void_p * 7)(*[w.data_ptr() for w in Wd])
            bp = (C.c_void_p * 7)(*[b.data_ptr() for b in bd])
            check(lib().ndp_planner_load(self._h, Wp, bp, 7, stream_ptr(self.device)))
            torch.cuda.current_stream(self.device).synchronize()   # sources may now be released

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                lib().ndp_planner_destroy(h)
            except Exception:
                pass
            self._h = None

    # ---- K1 ---------------------------------------------------------------------------------
    def rollout_score(self, e0, goal, noise, want_state=False):
        """e0, goal [G,E]; noise [T,G,K,nd] -> (cost[G,K], a0[G,K,4], state_T[G,K,E]|None).
        With precision "f16+refine" this is the single-pass bulk estimate (no refine)."""
        require_cuda(e0, goal, noise)
        e0c, gc, nz = f32c(e0), f32c(goal), f32c(noise)
        T, G, K, nd = nz.shape
        if T != self.T or nd != self.nd or tuple(e0c.shape) != (G, self.E) or tuple(gc.shape) != (G, self.E):
            raise _lib.NdpError(f"shape mismatch: noise {tuple(nz.shape)} e0 {tuple(e0c.shape)} for "
                                f"T={self.T} nd={self.nd} E={self.E}")
        dev = self.device
        cost = torch.empty(G, K, dtype=torch.float32, device=dev)
        a0 = torch.empty(G, K, 4, dtype=torch.float32, device=dev)
        sT = torch.empty(G, K, self.E, dtype=torch.float32, device=dev) if want_state else None
        n = C.c_size_t()
        check(lib().ndp_planner_workspace_bytes(self._h, G, K, C.byref(n)))
        ws = torch.empty(max(n.value, 8), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            check(lib().ndp_rollout_score(self._h, ptr(e0c), ptr(gc), ptr(nz), G, K, ptr(cost), ptr(a0), ptr(sT),
                                          ptr(ws), ws.numel(), stream_ptr(dev)))
        return cost, a0, sT

    # ---- K1 + K3 (+ refine) + gather: one C call for rollout and selection ------------------------------
    def plan(self, e0, goal, noise, topk: int = 0, idx_base: int = 0, want_state=False, eps: float = 0.0,
             refine_stream=None, fallback: bool = True) -> Plan:
        """mpc_eval.py:124-165.  precision "f16+refine": all K candidates are scored single-pass, the 128 best
        estimates are re-rolled with split operands and the plan is chosen from those fp32-grade costs;
        `Plan.verified[g]` tells whether that provably equals choosing from exact costs of all K.
        `refine_stream`: run the (single-tile, latency-bound) refine half on that stream so it overlaps the
        bulk half of the next plan; the returned Plan carries a `ready` event — call `plan.wait()` before
        reading it from another stream.
        `fallback` (f16+refine): goals whose refined selection cannot be proven exact are re-rolled over all K with
        split operands on the device (ndp_plan_fallback; `verified` 2 instead of 0)."""
        if want_state:
            if self.precision == "f16+refine":
                raise _lib.NdpError("want_state is not available with precision 'f16+refine'")
            cost, a0, sT = self.rollout_score(e0, goal, noise, True)
            packed, tk_idx, tk_cost = select_plan(cost, topk, idx_base)
            G, K = cost.shape
            verified = torch.ones(G, dtype=torch.int32, device=self.device)
        else:
            require_cuda(e0, goal, noise)
            e0c, gc, nz = f32c(e0), f32c(goal), f32c(noise)
            T, G, K, nd = nz.shape
            if T != self.T or nd != self.nd or tuple(e0c.shape) != (G, self.E) or tuple(gc.shape) != (G, self.E):
                raise _lib.NdpError(f"shape mismatch: noise {tuple(nz.shape)} e0 {tuple(e0c.shape)} for "
                                    f"T={self.T} nd={self.nd} E={self.E}")
            dev = self.device
            cost = torch.empty(G, K, dtype=torch.float32, device=dev)
            a0 = torch.empty(G, K, 4, dtype=torch.float32, device=dev)
            packed = torch.empty(G, dtype=torch.int64, device=dev)
            tk_idx = torch.empty(G, topk, dtype=torch.int32, device=dev) if topk > 0 else None
            tk_cost = torch.empty(G, topk, dtype=torch.float32, device=dev) if topk > 0 else None
            verified = torch.empty(G, dtype=torch.int32, device=dev)
            n = C.c_size_t()
            check(lib().ndp_plan_workspace_bytes(self._h, G, K, topk, C.byref(n)))
            ws = torch.empty(max(n.value, 256), dtype=torch.uint8, device=dev)
            args = (self._h, ptr(e0c), ptr(gc), ptr(nz), G, K, topk, idx_base, eps, ptr(cost), ptr(a0), ptr(packed),
                    ptr(tk_idx), ptr(tk_cost), ptr(verified), ptr(ws), ws.numel())
            sT = None
            best_action = torch.zeros(G, 4, dtype=torch.float32, device=dev)
            best_idx = torch.empty(G, dtype=torch.int32, device=dev)
            if refine_stream is not None and self.precision == "f16+refine":
                main = torch.cuda.current_stream(dev)
                with torch.cuda.device(dev):
                    check(lib().ndp_plan_select_bulk(*args, C.c_void_p(main.cuda_stream)))
                    refine_stream.wait_stream(main)
                    side = C.c_void_p(refine_stream.cuda_stream)
                    check(lib().ndp_plan_select_refine(*args, side))
                    if fallback:
                        check(lib().ndp_plan_fallback(*args, side))
                    check(lib().ndp_gather_best_action(ptr(a0), ptr(packed), G, K, idx_base, ptr(best_action),
                                                       ptr(best_idx), side))
                for t_ in (e0c, gc, nz, cost, a0, packed, tk_idx, tk_cost, verified, ws, best_action, best_idx):
                    if t_ is not None:
                        t_.record_stream(refine_stream)
                ready = torch.cuda.Event()
                ready.record(refine_stream)
                return Plan(best_idx, best_action, packed, cost, a0, tk_idx, tk_cost, sT, verified, ready,
                            (self._h, G, K, topk, ws) if self.precision == "f16+refine" else None)
            with torch.cuda.device(dev):
                check(lib().ndp_plan_select(*args, stream_ptr(dev)))
                if fallback and self.precision == "f16+refine":
                    check(lib().ndp_plan_fallback(*args, stream_ptr(dev)))
                check(lib().ndp_gather_best_action(ptr(a0), ptr(packed), G, K, idx_base, ptr(best_action),
                                                   ptr(best_idx), stream_ptr(dev)))
            return Plan(best_idx, best_action, packed, cost, a0, tk_idx, tk_cost, sT, verified, None,
                        (self._h, G, K, topk, ws) if self.precision == "f16+refine" else None)
        best_action = torch.zeros(G, 4, dtype=torch.float32, device=self.device)
        best_idx = torch.empty(G, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().ndp_gather_best_action(ptr(a0), ptr(packed), G, K, idx_base, ptr(best_action),
                                               ptr(best_idx), stream_ptr(self.device)))
        return Plan(best_idx, best_action, packed, cost, a0, tk_idx, tk_cost, sT, verified)

    def ring(self, G: int, K: int, topk: int = 0, depth: int = 4, n_streams: int = 2, idx_base: int = 0,
             comm=None, fallback: bool = False) -> "PlanRing":
        """pre-allocated plan slots + side streams for a steady planning loop (see PlanRing)"""
        return PlanRing(self, G, K, topk, depth, n_streams, idx_base, comm, fallback)

    # ---- single-call module semantics (drop-in for generator(z) / fwd_model(state, a)) -------
    def generator(self, z):
        """models/gan.py:79-86 — z [B, 2E+nd] -> actions [B,4]."""
        require_cuda(z)
        zc = f32c(z)
        out = torch.empty(zc.shape[0], 4, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().ndp_generator_forward(self._h, ptr(zc), zc.shape[0], ptr(out), stream_ptr(self.device)))
        return out

    def forward_model(self, state, action):
        """forward_encoder.py:105-114 (eval mode) — returns state + residual, same shape as state."""
        require_cuda(state, action)
        sc, ac = f32c(state).reshape(state.shape[0], -1), f32c(action)
        out = torch.empty_like(sc)
        with torch.cuda.device(self.device):
            check(lib().ndp_forward_step(self._h, ptr(sc), ptr(ac), sc.shape[0], ptr(out), stream_ptr(self.device)))
        return out.view(state.shape)


class _Slot:
    __slots__ = ("cost", "a0", "packed", "tk_idx", "tk_cost", "verified", "ws", "best_action", "best_idx", "ready",
                 "bulk_done", "inputs", "used", "tail", "gather", "plan", "rec", "rec_all", "exch", "head", "side")


class PlanRing:
    """Steady-state planning loop without per-plan allocations or host stalls.

    `depth` slots hold every output and workspace of a plan; plan i writes slot i % depth.  The bulk
    rollout runs on ring-owned stream i % 2, ordered after the caller's stream (so its inputs are ready): two
    consecutive rollouts may overlap, and the last, partially filled wave of 128-candidate tiles of plan i
    shares the SMs with the first wave of plan i+1 instead of leaving them idle (512 tiles on 148 SMs are
    3.46 waves).  Everything after the rollout (refine list selection, split-operand
    re-roll, selection, verification, action gather and — when `comm` is given — the sharded record
    exchange and merge) on side stream i % n_streams, so the latency-bound tails of two consecutive
    plans overlap each other and the next bulk rollouts.  A returned Plan is valid until `depth - 1`
    further plans have been issued on the ring; `plan.wait()` orders the current stream after it.
    Reusing a slot makes the caller's stream wait (on the device) for the slot's previous tail."""

    def __init__(self, planner: FusedPlanner, G: int, K: int, topk: int = 0, depth: int = 4, n_streams: int = 2,
                 idx_base: int = 0, comm=None, fallback: bool = False):
        dev = planner.device
        # fallback=True appends ndp_plan_fallback to every tail.  Off by default in the pipelined loop: its (instantly
        # exiting) 213 KB-smem thread blocks can only be placed on SMs a bulk rollout has vacated, which stalls the
        # tail and steals slots from the next rollout; read `plan.verified` and call `ring.fallback(plan)` instead.
        self.fallback_always = fallback
        self.planner, self.G, self.K, self.topk, self.depth, self.idx_base, self.comm = \
            planner, G, K, topk, depth, idx_base, comm
        self.count = 0
        self.vlog = None                              # optional [n,G] int32 log of every plan's `verified` (log_verified)
        self.streams = [torch.cuda.Stream(device=dev, priority=-1) for _ in range(n_streams)]
        self.bulk_streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
        self.inputs_ready = [torch.cuda.Event() for _ in range(depth)]
        n = C.c_size_t()
        check(lib().ndp_plan_workspace_bytes(planner._h, G, K, topk, C.byref(n)))
        world = comm.world if comm is not None else 1
        if world * topk > 1024:
            raise _lib.NdpError("sharded top-k needs world*topk <= 1024")
        if comm is not None and depth > 7:
            raise _lib.NdpError("a sharded ring keeps depth + 1 record exchanges in flight: depth <= 7")
        self.slots = []
        f32, i32, i64 = torch.float32, torch.int32, torch.int64
        for _ in range(depth):
            s = _Slot()
            s.cost = torch.empty(G, K, dtype=f32, device=dev)
            s.a0 = torch.empty(G, K, 4, dtype=f32, device=dev)
            s.packed = torch.empty(G, dtype=i64, device=dev)
            s.tk_idx = torch.empty(G, topk, dtype=i32, device=dev) if topk else None
            s.tk_cost = torch.empty(G, topk, dtype=f32, device=dev) if topk else None
            s.verified = torch.empty(G, dtype=i32, device=dev)
            s.ws = torch.empty(max(n.value, 256), dtype=torch.uint8, device=dev)
            s.best_action = torch.zeros(G, 4, dtype=f32, device=dev)
            s.best_idx = torch.empty(G, dtype=i32, device=dev)
            s.ready, s.bulk_done, s.inputs, s.used = torch.cuda.Event(), torch.cuda.Event(), None, False
            s.tail = (G, K, topk, idx_base, 0.0, ptr(s.cost), ptr(s.a0), ptr(s.packed), ptr(s.tk_idx), ptr(s.tk_cost),
                      ptr(s.verified), ptr(s.ws), s.ws.numel())
            s.gather = (ptr(s.a0), ptr(s.packed), G, K, idx_base, ptr(s.best_action), ptr(s.best_idx))
            if comm is None:
                s.plan = Plan(s.best_idx, s.best_action, s.packed, s.cost, s.a0, s.tk_idx, s.tk_cost, None, s.verified,
                              s.ready, (planner._h, G, K, topk, s.ws) if planner.precision == "f16+refine" else None)
            else:
                W = 4 + topk
                s.rec = torch.empty(G, W, dtype=i64, device=dev)
                s.rec_all = torch.empty(world, G, W, dtype=i64, device=dev)
                g_packed, g_idx = torch.empty(G, dtype=i64, device=dev), torch.empty(G, dtype=i32, device=dev)
                g_act, g_ver = torch.empty(G, 4, dtype=f32, device=dev), torch.empty(G, dtype=i32, device=dev)
                g_tki = torch.empty(G, topk, dtype=i32, device=dev) if topk else None
                g_tkc = torch.empty(G, topk, dtype=f32, device=dev) if topk else None
                # ndp_shard_exchange: pack -> exchange (peer window over NVLink, or ncclAllGather) -> merge
                s.exch = (ptr(s.packed), ptr(s.best_action), ptr(s.verified), ptr(s.tk_idx), ptr(s.tk_cost), G, topk,
                          ptr(s.rec), ptr(s.rec_all), ptr(g_packed), ptr(g_idx), ptr(g_act), ptr(g_tki), ptr(g_tkc),
                          ptr(g_ver))
                s.plan = Plan(g_idx, g_act, g_packed, s.cost, s.a0, g_tki, g_tkc, None, g_ver, s.ready)
            self.slots.append(s)

    def log_verified(self, n: int) -> torch.Tensor:
        """keep the `verified` flags of the next n plans: plan i's tail copies its [G] flags into row i % n of the
        returned device tensor (one 4*G-byte device copy on the tail stream, no host synchronisation), so a pipelined
        loop can account for every plan after the fact although slots are reused"""
        self.vlog = torch.full((int(n), self.G), -1, dtype=torch.int32, device=self.planner.device)
        self.vlog_base = self.count
        return self.vlog

    def plan(self, e0, goal, noise) -> Plan:
        """mpc_eval.py:124-165 for e0, goal [G,E] and noise [T,G,K,nd]; returns the slot's Plan (see class doc)"""
        pl = self.planner
        require_cuda(e0, goal, noise)
        e0c, gc, nz = f32c(e0), f32c(goal), f32c(noise)
        if tuple(nz.shape) != (pl.T, self.G, self.K, pl.nd) or tuple(e0c.shape) != (self.G, pl.E) \
                or tuple(gc.shape) != (self.G, pl.E):
            raise _lib.NdpError(f"shape mismatch: noise {tuple(nz.shape)} e0 {tuple(e0c.shape)} for ring "
                                f"T={pl.T} G={self.G} K={self.K} nd={pl.nd} E={pl.E}")
        i = self.count
        self.count = i + 1
        s = self.slots[i % self.depth]
        side = self.streams[i % len(self.streams)]
        dev = pl.device
        main = torch.cuda.current_stream(dev)
        bulk = self.bulk_streams[i & 1]
        if s.used:
            main.wait_event(s.ready)          # the slot's previous tail is done: its inputs may be released ...
            bulk.wait_event(s.ready)          # ... and its outputs overwritten
        s.inputs, s.used = (e0c, gc, nz), True   # keep the inputs alive until the slot is reused
        ir = self.inputs_ready[i % self.depth]
        ir.record(main)
        bulk.wait_event(ir)
        L = lib()
        head = (pl._h, e0c.data_ptr(), gc.data_ptr(), nz.data_ptr())
        sp = C.c_void_p(side.cuda_stream)
        s.head, s.side = head, side
        with torch.cuda.device(dev):
            check(L.ndp_plan_select_bulk(*head, *s.tail, C.c_void_p(bulk.cuda_stream)))
            s.bulk_done.record(bulk)
            side.wait_event(s.bulk_done)
            check(L.ndp_plan_select_refine(*head, *s.tail, sp))
            if self.fallback_always and pl.precision == "f16+refine":
                check(L.ndp_plan_fallback(*head, *s.tail, sp))
            check(L.ndp_gather_best_action(*s.gather, sp))
            if self.comm is not None:
                check(L.ndp_shard_exchange(self.comm._h, *s.exch, sp))
            if self.vlog is not None:
                with torch.cuda.stream(side):
                    self.vlog[(i - self.vlog_base) % self.vlog.shape[0]].copy_(s.plan.verified, non_blocking=True)
            s.ready.record(side)
        return s.plan

    def fallback(self, plan: Plan) -> Plan:
        """Consumer-initiated exact path for a plan of this ring whose `verified` holds zeros (f16+refine): re-rolls the
        unverified goals over all K with split operands on the plan's tail stream (ndp_plan_fallback), gathers the
        action again and re-records `plan.ready`.  Call before 3 further plans have been issued."""
        if self.comm is not None:
            raise _lib.NdpError("ring.fallback is not available on a sharded ring; build it with fallback=True")
        for s in self.slots:
            if s.plan is plan and s.used:
                sp = C.c_void_p(s.side.cuda_stream)
                with torch.cuda.device(self.planner.device):
                    check(lib().ndp_plan_fallback(*s.head, *s.tail, sp))
                    check(lib().ndp_gather_best_action(*s.gather, sp))
                    s.ready.record(s.side)
                return plan
        raise _lib.NdpError("plan does not belong to this ring (or its slot has been reused)")
