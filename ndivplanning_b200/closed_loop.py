"""Batched receding-horizon sweep — the outer ``image_num`` loop of the reference
(mpc_eval.py:112-183; complete_eval.py:111-156 runs the same loop over many trajectories) for G
independent (start, goal) pairs at once, entirely stream-ordered on the device:

    for s in range(steps):                                  # mpc_eval.py:112
        T_eff = min(Th, steps - s)                          # truncated horizon near the end (:127)
        plan_s = K candidates x T_eff steps per goal        # :124-165  (FusedPlanner, K1 + K3)
        cur    = fwd_model(cur, best_action_s)              # :167-169  (ndp_forward_step)
        err_s  = mse(cur, state_fut_s)                      # :173-176

No host synchronisation happens inside the loop: noise comes from the device generator
(``torch.rand`` on the planner's GPU — the sampler of mpc_eval.py:78-83 moved next to its consumer) or
from a caller-supplied ``noise_fn`` (tests pass the oracle's CPU draws), the winning action is gathered
on the device and fed to the forward model by pointer.  Goals shard across GPUs with no collective
(BASELINE config 5): every rank runs this sweep on its G/P goals.
"""
from dataclasses import dataclass
from typing import Callable, Optional

import torch

from . import _lib
from .planner import FusedPlanner


@dataclass
class SweepResult:
    actions: torch.Tensor        # [S,G,4]  chosen first actions (best_action_list, mpc_eval.py:171)
    best_idx: torch.Tensor       # [S,G]    int32 winning candidate per outer step
    states: torch.Tensor         # [S,G,E]  model-predicted state after applying the chosen action
    image_err: Optional[torch.Tensor]    # [S,G] mse(cur, state_fut) per outer step (:173) or None
    action_err: Optional[torch.Tensor]   # [G]   mse(actions_ref, actions) per trajectory (:178-181) or None
    verified: torch.Tensor       # [S,G]    int32 (f16+refine: selection proven exact)
    candidate_steps: int         # G * K * sum(T_eff): the work this sweep scored


class ClosedLoopSweep:
    def __init__(self, generator, fwd_model, *, horizon: int, rollouts: int, precision: str = "f16+refine",
                 device=None):
        self.gen, self.fwd, self.Th, self.K, self.precision = generator, fwd_model, int(horizon), int(rollouts), precision
        self._planners = {}
        self._dev = device
        p = self.planner(self.Th)
        self.device, self.nd, self.E = p.device, p.nd, p.E

    def planner(self, T: int) -> FusedPlanner:
        """one planner handle per effective horizon (weights are packed once per handle)"""
        if T not in self._planners:
            self._planners[T] = FusedPlanner(self.gen, self.fwd, horizon=T, precision=self.precision, device=self._dev)
        return self._planners[T]

    @torch.no_grad()
    def run(self, start: torch.Tensor, goal: torch.Tensor, steps: int, state_fut: Optional[torch.Tensor] = None,
            actions_ref: Optional[torch.Tensor] = None, noise_fn: Optional[Callable[[int, int], torch.Tensor]] = None,
            generator: Optional[torch.Generator] = None, receding: bool = True) -> SweepResult:
        """start, goal [G,E]; state_fut [S,G,E] (targets of the image error, optional); actions_ref [S,G,4]
        (ground-truth actions, optional); noise_fn(step, T_eff) -> [T_eff,G,K,nd] overrides the device sampler.
        receding=True truncates the horizon to the steps left (mpc_eval.py:127); False always plans Th steps
        (MPC_gym_eval.py:188)."""
        _lib.require_cuda(start, goal)
        dev = self.device
        G = start.shape[0]
        cur = _lib.f32c(start).reshape(G, self.E).to(dev)
        goal = _lib.f32c(goal).reshape(G, self.E).to(dev)
        S = int(steps)
        actions = torch.empty(S, G, 4, dtype=torch.float32, device=dev)
        best_idx = torch.empty(S, G, dtype=torch.int32, device=dev)
        states = torch.empty(S, G, self.E, dtype=torch.float32, device=dev)
        verified = torch.empty(S, G, dtype=torch.int32, device=dev)
        image_err = torch.empty(S, G, dtype=torch.float32, device=dev) if state_fut is not None else None
        work = 0
        for s in range(S):
            T_eff = min(self.Th, S - s) if receding else self.Th
            pl = self.planner(T_eff)
            if noise_fn is not None:
                noise = noise_fn(s, T_eff).to(dev, non_blocking=True)
            else:
                noise = torch.rand(T_eff, G, self.K, self.nd, device=dev, generator=generator)   # U[0,1), :80
            plan = pl.plan(cur, goal, noise)
            cur = pl.forward_model(cur, plan.best_action)                                        # :167-169
            actions[s].copy_(plan.best_action)
            best_idx[s].copy_(plan.best_idx)
            verified[s].copy_(plan.verified)
            states[s].copy_(cur)
            if image_err is not None:
                image_err[s] = ((cur - state_fut[s].to(dev).reshape(G, self.E)) ** 2).mean(dim=1)  # :173
            work += G * self.K * T_eff
        action_err = None
        if actions_ref is not None:
            action_err = ((actions_ref.to(dev) - actions) ** 2).mean(dim=(0, 2))                  # :178-181
        return SweepResult(actions, best_idx, states, image_err, action_err, verified, work)

    def capture(self, start: torch.Tensor, goal: torch.Tensor, steps: int, state_fut: Optional[torch.Tensor] = None,
                actions_ref: Optional[torch.Tensor] = None, noise_fn: Optional[Callable[[int, int], torch.Tensor]] = None,
                receding: bool = True):
        """The whole closed loop — `steps` plans, forward-model steps and error accumulations — as ONE CUDA graph
        (SURVEY.md 8f N1).  `start`, `goal` (and `state_fut`, `actions_ref`, the tensors `noise_fn` returns) are the
        graph's static inputs: refill them in place and `graph.replay()`; the returned SweepResult holds the static
        outputs.  Without `noise_fn` the noise comes from the device generator inside the graph (a fresh draw per
        replay: torch advances the Philox offset of captured generators).  One warm-up run is executed eagerly."""
        dev = self.device
        _lib.require_cuda(start, goal)
        S = int(steps)
        for s_ in range(S):                                   # weights are packed outside the capture
            self.planner(min(self.Th, S - s_) if receding else self.Th)
        self.run(start, goal, S, state_fut, actions_ref, noise_fn, None, receding)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                res = self.run(start, goal, S, state_fut, actions_ref, noise_fn, None, receding)
        torch.cuda.current_stream(dev).wait_stream(side)
        return g, res
