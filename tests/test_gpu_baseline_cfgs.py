"""GPU: parity at the sizes BASELINE.json names (cfg2 / cfg3 / cfg4-shard / cfg5-shard) against the CPU oracle, and
the soundness of the `verified` flag of the headline mode ("f16+refine") in regimes chosen to break it.

Oracle: fp64 twin of oracle/ndiv_oracle.py (arbiter) + fp32 port; ndiv at k=4096 through the row-blocked fp64
evaluation (the reference's own [k,k,C] difference tensor would need 4.3 GB).  Bars: argmin index and top-k set
bit-exact (ambiguity protocol of test_gpu_mpc.py where the fp64 gap is inside the mode's error bound), refined costs
and the ndiv loss within 1e-4 relative.
"""
import numpy as np
import pytest
import torch

from oracle import ndiv_oracle as O
from ndivplanning_b200 import synthetic
from test_gpu_mpc import COST_RTOL, _check_selection_bound, _oracle_models
from test_gpu_ndiv import RTOL, _grad_close_arbiter

pytestmark = pytest.mark.gpu


def _true_dev(cost_plan, c64):
    """largest |sqrt(cost) - sqrt(exact)| over all candidates (the quantity the verification bound eta must cover)"""
    return float(np.abs(np.sqrt(cost_plan.astype(np.float64)) - np.sqrt(c64)).max())


@pytest.mark.parametrize("name,K,T", [("cfg2", 65536, 8), ("cfg4_shard", 131072, 16)])
def test_baseline_mpc_f16refine_vs_oracle(cuda, name, K, T):
    """BASELINE cfg2 (K=65,536, T=8) and one rank's shard of cfg4 (K=1,048,576/8, T=16) in the headline mode against
    the full-K fp64 oracle (and the fp32 port): plan verified, argmin and top-16 set equal, refined costs within 1e-4,
    bulk estimates inside the bound the verification assumed; the pipelined PlanRing returns the same plan."""
    from ndivplanning_b200 import FusedPlanner
    nd, E, H, topk = 2, 128, 1024, 16
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(1, K, T, nd, E, seed=5)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision="f16+refine")
    d = [t.cuda() for t in (e0, goal, noise)]
    plan = pl.plan(*d, topk=topk, fallback=False)
    info = plan.verify_info().cpu().numpy()
    g64, f64 = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    ref64 = O.mpc_plan(g64, f64, e0[0].double(), goal[0].double(), noise[:, 0].double(), topk=topk)
    c64 = ref64["cost"].numpy()
    assert plan.verified.cpu().tolist() == [1], info
    best, tk = int(plan.best_idx[0]), plan.topk_idx[0].cpu().tolist()
    _check_selection_bound(c64, best, tk, topk, 2e-5)
    order = np.argsort(c64, kind="stable")
    assert best == int(order[0]) and set(tk) == set(order[:topk].tolist())      # sep30 weights: unambiguous
    c = plan.cost[0].cpu().numpy()
    for i in set(tk + [best]):
        assert abs(c[i] - c64[i]) <= COST_RTOL * c64[i]
    np.testing.assert_allclose(plan.topk_cost[0].cpu().numpy(), c64[tk], rtol=COST_RTOL)
    np.testing.assert_allclose(c, c64, rtol=1.2e-3)                               # bulk: single-pass estimates
    assert _true_dev(c, c64) <= info[0, 1], (_true_dev(c, c64), info)             # the assumed bound covers ALL K
    assert info[0, 0] > 0
    np.testing.assert_allclose(plan.best_action[0].cpu().numpy(), ref64["a0"][best].numpy(), rtol=1e-4, atol=1e-5)
    if name == "cfg2":                                                            # fp32 port (the reference's dtype)
        gen, fwd = _oracle_models(gen_p, fwd_p, nd, E, H)
        ref = O.mpc_plan(gen, fwd, e0[0], goal[0], noise[:, 0], topk=topk)
        assert best == ref["best_idx"] and set(tk) == set(ref["topk_idx"].tolist())
        np.testing.assert_allclose(plan.topk_cost[0].cpu().numpy(), ref["cost"][tk].numpy(), rtol=COST_RTOL)
    ring = pl.ring(1, K, topk)
    for _ in range(3):
        rp = ring.plan(*d)
    rp.wait()
    torch.cuda.synchronize()
    for a, b in ((rp.best_idx, plan.best_idx), (rp.topk_idx, plan.topk_idx), (rp.topk_cost, plan.topk_cost),
                 (rp.verified, plan.verified), (rp.best_action, plan.best_action), (rp.cost, plan.cost)):
        assert torch.equal(a, b)


def test_baseline_cfg5_shard_closed_loop_vs_oracle(cuda):
    """a rank's shard of BASELINE cfg5 (goals sharded, K=4,096, T=8, 7 receding outer steps; G=32 of the 512 goals a
    rank owns): every outer step's selection is checked against the fp64 oracle plan computed FROM THE SWEEP'S OWN
    current state (so one ambiguous step cannot derail the comparison), the state update against the oracle forward
    model, and every plan must end verified (1) or exactly re-rolled (2)."""
    from ndivplanning_b200.closed_loop import ClosedLoopSweep
    G, K, Th, S, nd, E, H = 32, 4096, 8, 7, 2, 128, 1024
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H)
    g64, f64 = _oracle_models(gen, fwd, nd, E, H, double=True)
    torch.manual_seed(21)
    start, goal = torch.randn(G, E), torch.randn(G, E)
    noises = [torch.rand(min(Th, S - s), G, K, nd) for s in range(S)]
    sw = ClosedLoopSweep(gen, fwd, horizon=Th, rollouts=K, precision="f16+refine")
    res = sw.run(start.cuda(), goal.cuda(), S, noise_fn=lambda s, T: noises[s])
    assert res.candidate_steps == G * K * sum(min(Th, S - s) for s in range(S))
    ver = res.verified.cpu()
    assert int(ver.min()) >= 1 and int(ver.max()) <= 2
    states, actions, best = res.states.cpu(), res.actions.cpu(), res.best_idx.cpu()
    n_exact = 0
    for g in range(G):
        for s in range(S):
            cur = start[g] if s == 0 else states[s - 1, g]
            ref = O.mpc_plan(g64, f64, cur.double(), goal[g].double(), noises[s][:, g].double())
            c64 = ref["cost"].numpy()
            b = int(best[s, g])
            _check_selection_bound(c64, b, [], 0, 2e-5)
            n_exact += int(b == ref["best_idx"])
            np.testing.assert_allclose(actions[s, g].numpy(), ref["a0"][b].numpy(), rtol=1e-4, atol=1e-5)
            with torch.no_grad():
                nxt = f64(cur.double().view(1, E, 1, 1), ref["a0"][b][None]).view(E)
            np.testing.assert_allclose(states[s, g].numpy(), nxt.numpy(), rtol=1e-4, atol=1e-5)
    assert n_exact >= G * S - 2, n_exact          # ambiguous (gap < 1.6e-4 relative) steps are rare


def test_baseline_cfg3_ndiv_vs_chunked_fp64(cuda):
    """BASELINE cfg3: N=1, k=4096, Dx=32, Dz=16 forward + backward (the tensor-core Gram path) against the row-blocked
    fp64 oracle; hinge-boundary pairs get the fp64-arbiter slack of test_gpu_ndiv.py"""
    from ndivplanning_b200 import diversity as div
    N, k, Dx, Dz = 1, 4096, 32, 16
    x, z = synthetic.make_ndiv_inputs(N, k, Dx, Dz, seed=2)
    xg = x.cuda().requires_grad_(True)
    loss = div.compute_pairwise_divergence(xg, z.cuda())
    loss.backward()
    l64, g64, sx64, sz64 = O.ndiv_fwd_bwd_chunked(x[0].double(), z[0].double(), block=512)
    np.testing.assert_allclose(loss.item(), l64.item(), rtol=RTOL)
    amb = _grad_close_arbiter(xg.grad.cpu(), x, z, g64[None].numpy())
    assert amb < 2048, amb          # hinge-boundary pairs among 16.8 M (fp32 round-off of the normalised distances)
    sx, sz = div.ndiv_rowsums(x.cuda(), z.cuda())
    np.testing.assert_allclose(sx[0].cpu().numpy(), sx64.numpy(), rtol=1e-5)
    np.testing.assert_allclose(sz[0].cpu().numpy(), sz64.numpy(), rtol=1e-5)


# ---- soundness of `verified`: regimes chosen to stress the error model of the single-pass estimate ---------------
def _regime(name):
    nd, E = 2, 128
    if name == "H2048_nd15":          # the shape whose single-pass cost error reaches 1.2e-3 (test_gpu_mpc.py)
        G, K, T, H, nd, sep, scale = 2, 2048, 5, 2048, 15, 30.0, 1.0
    elif name == "long_horizon":
        G, K, T, H, sep, scale = 2, 2048, 32, 256, 30.0, 1.0
    elif name == "weak_separation":   # short horizon, no sep scaling: costs closer than the estimate's error
        G, K, T, H, sep, scale = 6, 1024, 1, 256, 1.0, 1.0
    elif name == "scaled_dynamics":   # forward model fc2 x1.25: round-off is amplified step by step
        G, K, T, H, sep, scale = 2, 2048, 6, 512, 30.0, 1.25
    elif name == "near_goal":         # goal = terminal state of one candidate: cost -> 0, relative error unbounded
        G, K, T, H, sep, scale = 2, 2048, 4, 256, 30.0, 1.0
    else:
        raise KeyError(name)
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H, seed=1, sep=sep)
    if scale != 1.0:
        with torch.no_grad():
            fwd.fc2.weight.mul_(scale)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, nd, E, seed=17)
    return gen, fwd, e0, goal, noise, T


@pytest.mark.parametrize("name", ["H2048_nd15", "long_horizon", "weak_separation", "scaled_dynamics", "near_goal"])
def test_verified_flag_is_sound(cuda, name):
    """verified == 1 must imply that the plan equals the selection from split-operand (fp32-grade) costs over ALL K
    candidates, and the error bound the proof assumed (eta) must cover the true error of every estimate; goals where
    it cannot be shown come back 0 (fallback off) and are re-rolled to 2 by the fallback — never a wrong 1."""
    from ndivplanning_b200 import FusedPlanner
    gen, fwd, e0, goal, noise, T = _regime(name)
    topk = 8
    d = [t.cuda() for t in (e0, goal, noise)]
    exact_pl = FusedPlanner(gen, fwd, horizon=T, precision="f16x3")
    if name == "near_goal":
        _, _, sT = FusedPlanner(gen, fwd, horizon=T, precision="fp32").rollout_score(*d, want_state=True)
        goal = sT[:, 777].clone().cpu() + 1e-6 * torch.randn(goal.shape)
        d[1] = goal.cuda()
    exact, _, _ = exact_pl.rollout_score(*d)
    est, _, _ = FusedPlanner(gen, fwd, horizon=T, precision="f16").rollout_score(*d)
    pl = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine")
    plan = pl.plan(*d, topk=topk, fallback=False)
    info = plan.verify_info().cpu().numpy()
    ver = plan.verified.cpu().tolist()
    ex, es = exact.cpu().numpy().astype(np.float64), est.cpu().numpy().astype(np.float64)
    G = ex.shape[0]
    for g in range(G):
        if not (np.isfinite(es[g]).all() and np.isfinite(ex[g]).all()):   # overflowed fp16 range: never a proof
            assert ver[g] == 0
            continue
        true_dev = float(np.abs(np.sqrt(es[g]) - np.sqrt(ex[g])).max())
        assert info[g, 1] >= true_dev, (name, g, info[g], true_dev)        # eta covers every candidate's error
        assert (info[g, 0] > 0) == (ver[g] == 1)
        if ver[g] == 1:
            order = np.argsort(ex[g], kind="stable")
            assert int(plan.best_idx[g]) == int(order[0]), (name, g)
            assert set(plan.topk_idx[g].cpu().tolist()) == set(order[:topk].tolist()), (name, g)
    if name in ("H2048_nd15", "long_horizon", "near_goal"):
        assert 1 in ver, (name, ver, info)                                  # the proof is not vacuous
    if name == "weak_separation":
        assert 0 in ver, (name, ver, info)
    # with the fallback every goal ends exact: 1 (proven) or 2 (re-rolled == the f16x3 planner's plan)
    full = exact_pl.plan(*d, topk=topk)
    fb = pl.plan(*d, topk=topk)
    v2 = fb.verified.cpu().tolist()
    assert set(v2) <= {1, 2} and [v == 1 for v in v2] == [v == 1 for v in ver]
    for g in range(G):
        assert int(fb.best_idx[g]) == int(full.best_idx[g])
        assert set(fb.topk_idx[g].cpu().tolist()) == set(full.topk_idx[g].cpu().tolist())
        if v2[g] == 2:
            assert torch.equal(fb.cost[g], full.cost[g]) and torch.equal(fb.topk_idx[g], full.topk_idx[g])


def test_caller_eps_only_tightens(cuda):
    """eps > 0 is a floor on the assumed relative cost error: it can turn a 1 into a 0, never a 0 into a 1"""
    from ndivplanning_b200 import FusedPlanner
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256)
    e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(4, 1024, 3, 2, 128, seed=8))
    pl = FusedPlanner(gen, fwd, horizon=3, precision="f16+refine")
    base = pl.plan(e0, goal, noise, topk=4, fallback=False).verified.cpu().tolist()
    huge = pl.plan(e0, goal, noise, topk=4, fallback=False, eps=0.9).verified.cpu().tolist()
    assert huge == [0] * 4
    mid = pl.plan(e0, goal, noise, topk=4, fallback=False, eps=1e-5).verified.cpu().tolist()
    assert all(m <= b for m, b in zip(mid, base))
