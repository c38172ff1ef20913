"""GPU: K1 fused rollout (+K3 selection) against the oracle loop (mpc_eval.py:124-165) on the
same seeded weights/inputs, and against the golden vectors of the unmodified reference.

Tolerances (north_star): fp32 costs within 1e-4 relative; argmin index and top-k set
bit-exact.  fp32 summation order differs between the kernel and MKL, so index equality is
asserted where the fp64 arbiter's gap exceeds the fp32 error bound (class 1) and otherwise the
chosen candidate must be fp64-optimal to 1e-6 (class 2) — SURVEY.md §7.4-1c.  On the sep30
bench weights every case must be class 1.
"""
import numpy as np
import pytest
import torch

from oracle import ndiv_oracle as O
from ndivplanning_b200 import synthetic

pytestmark = pytest.mark.gpu
COST_RTOL = 1e-4


def _oracle_models(gen_p, fwd_p, nd, E, H, double=False):
    gen, fwd = O.GeneratorMLP(nd, cond_dim=2 * E).eval(), O.LatentForward(E, H).eval()
    gen.load_state_dict(gen_p.state_dict()); fwd.load_state_dict(fwd_p.state_dict())
    return (gen.double(), fwd.double()) if double else (gen, fwd)


def _check_selection(cost_gpu, best_gpu, topk_gpu, cost64, topk):
    """class-1/class-2 protocol; returns 1 or 2"""
    order = np.argsort(cost64, kind="stable")
    gap_rel = (cost64[order[1]] - cost64[order[0]]) / cost64[order[0]] if len(order) > 1 else 1.0
    klass = 1 if gap_rel > 8 * 3e-7 else 2
    if klass == 1:
        assert best_gpu == order[0]
    else:
        assert cost64[best_gpu] <= cost64[order[0]] * (1 + 1e-6)
    if topk and len(order) > topk:
        bgap = (cost64[order[topk]] - cost64[order[topk - 1]]) / cost64[order[topk - 1]]
        if bgap > 8 * 3e-7:
            assert set(topk_gpu) == set(order[:topk].tolist())
        else:
            assert max(cost64[i] for i in topk_gpu) <= cost64[order[topk - 1]] * (1 + 1e-6)
    return klass


# E = 256: the alternative reading of BASELINE cfg2's "embedding dim 256" (SURVEY.md 8a note) — fp32 kernel, 64-candidate
# tiles; K = 65 / 130 / 257 are ragged against that tile, H = 200 is padded to 256 inside the library
@pytest.mark.parametrize("G,K,T,H,nd,E", [(1, 64, 1, 128, 2, 128), (1, 128, 3, 256, 2, 128), (2, 300, 4, 256, 2, 128),
                                          (1, 1024, 8, 1024, 2, 128), (3, 130, 2, 384, 5, 128), (1, 257, 5, 128, 1, 128),
                                          (1, 65, 1, 128, 2, 256), (2, 130, 3, 256, 2, 256), (1, 257, 4, 200, 5, 256),
                                          (1, 1024, 8, 1024, 2, 256)])
def test_rollout_matches_oracle(cuda, G, K, T, H, nd, E):
    from ndivplanning_b200 import FusedPlanner
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, nd, E, seed=K + T)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision="fp32")
    plan = pl.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=min(8, K), want_state=True)
    gen, fwd = _oracle_models(gen_p, fwd_p, nd, E, H)
    g64, f64 = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    for g in range(G):
        ref = O.mpc_plan(gen, fwd, e0[g], goal[g], noise[:, g], topk=min(8, K))
        ref64 = O.mpc_plan(g64, f64, e0[g].double(), goal[g].double(), noise[:, g].double())
        c = plan.cost[g].cpu().numpy()
        np.testing.assert_allclose(c, ref["cost"].numpy(), rtol=COST_RTOL)
        np.testing.assert_allclose(c, ref64["cost"].numpy(), rtol=COST_RTOL)
        np.testing.assert_allclose(plan.a0[g].cpu().numpy(), ref["a0"].numpy(), rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(plan.state_T[g].cpu().numpy(), ref64["state_T"].numpy(), rtol=1e-3, atol=1e-4)
        klass = _check_selection(c, int(plan.best_idx[g]), plan.topk_idx[g].cpu().tolist(),
                                 ref64["cost"].numpy(), min(8, K))
        assert klass == 1, "sep30 weights must separate candidates (SURVEY §7.4-1a)"
        # selected action = step-0 action of the winner (mpc_eval.py:150-151,165)
        np.testing.assert_allclose(plan.best_action[g].cpu().numpy(), ref["a0"][int(plan.best_idx[g])].numpy(),
                                   rtol=1e-4, atol=1e-5)
        # the kernel's own argmin is exactly the first-min of the kernel's own costs
        assert int(plan.best_idx[g]) == O.select_first_min(plan.cost[g].cpu())


@pytest.mark.parametrize("prec", ["fp32", None])
@pytest.mark.parametrize("name", ["small", "cfg1"])
def test_plan_golden(cuda, golden, name, prec):
    """explicit plan from the reference's own Decoder (golden), K=64 / K=1024 (BASELINE config 1); prec=None is the
    planner's default (f16+refine with the exact fallback): same selection, fp32-grade costs on the selected entries"""
    from ndivplanning_b200 import FusedPlanner
    g = golden["mpc"]
    K, Th, seq, n_traj, E, H, nd, seed = [int(v) for v in g[f"{name}_params"]]
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(1, K, Th, nd, E, seed=21)
    pl = FusedPlanner(gen_p, fwd_p, horizon=Th, precision=prec)
    plan = pl.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=8)
    if prec is None:
        assert pl.precision == "f16+refine" and plan.verified.cpu().tolist()[0] in (1, 2)
        sel = plan.topk_idx[0].cpu().tolist()
        np.testing.assert_allclose(plan.cost[0].cpu().numpy()[sel], g[f"{name}_plan_cost"][sel], rtol=COST_RTOL)
        np.testing.assert_allclose(plan.a0[0].cpu().numpy()[sel], g[f"{name}_plan_a0"][sel], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(plan.cost[0].cpu().numpy(), g[f"{name}_plan_cost"], rtol=1.2e-3)
    else:
        np.testing.assert_allclose(plan.cost[0].cpu().numpy(), g[f"{name}_plan_cost"], rtol=COST_RTOL)
        np.testing.assert_allclose(plan.a0[0].cpu().numpy(), g[f"{name}_plan_a0"], rtol=1e-4, atol=1e-5)
    assert _check_selection(plan.cost[0].cpu().numpy(), int(plan.best_idx[0]), plan.topk_idx[0].cpu().tolist(),
                            g[f"{name}_plan_cost64"], 8) == 1
    assert int(plan.best_idx[0]) == int(g[f"{name}_plan_best"])
    assert set(plan.topk_idx[0].cpu().tolist()) == set(g[f"{name}_plan_topk"].tolist())


@pytest.mark.parametrize("name", ["small", "cfg1"])
def test_fetch_push_control_evaluation_golden(cuda, golden, name):
    """the drop-in function reproduces the unmodified reference's return tuple"""
    from torch import nn
    from ndivplanning_b200 import fetch_push_control_evaluation
    from oracle.ref_stubs import DotMap
    g = golden["mpc"]
    K, Th, seq, n_traj, E, H, nd, seed = [int(v) for v in g[f"{name}_params"]]
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    images, actions = torch.from_numpy(g[f"{name}_images"]), torch.from_numpy(g[f"{name}_actions"])

    class DS(torch.utils.data.Dataset):
        seq_length = seq
        def __len__(self): return n_traj
        def __getitem__(self, i):
            return images[i][:, :, None, None], torch.zeros(seq, 25), actions[i], torch.zeros(3)

    cfg = DotMap(random_seed=seed, gpu_id=0, log_port=0,
                 evaluation=dict(num_sample=1, noise_dim=nd, batch_size=1, threshold=0.05),
                 mpc=dict(rollouts=K, time_horizon=Th))
    a, i, best = fetch_push_control_evaluation(nn.Identity(), fwd_p, gen_p, DS(), cfg, return_indices=True)
    assert np.array_equal(best.numpy(), g[f"{name}_best"])
    np.testing.assert_allclose([a, i], g[f"{name}_tuple"], rtol=1e-4)


def test_module_semantics(cuda, golden):
    """generator(z) and fwd_model(state, a) single-call entry points (gan.py:79-86, forward_encoder.py:105-114)"""
    from ndivplanning_b200 import FusedPlanner
    g = golden["gen"]
    torch.manual_seed(0)
    gen = synthetic.GeneratorMLP(2)
    fwd = synthetic.LatentForward(128, 256).eval()
    pl = FusedPlanner(gen, fwd, horizon=1)
    a = pl.generator(torch.from_numpy(g["nd2_z"]).cuda())
    np.testing.assert_allclose(a.cpu().numpy(), g["nd2_a"], rtol=1e-4, atol=1e-6)
    s, act = torch.randn(9, 128, 1, 1), torch.randn(9, 4)
    with torch.no_grad():
        want = fwd(s, act)
    got = pl.forward_model(s.cuda(), act.cuda())
    assert got.shape == s.shape
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-4, atol=1e-5)


def test_full_size_properties(cuda):
    """BASELINE config 2 size (K=65,536, T=8): size-independent checks — permutation of
    candidates permutes costs bit-exactly; duplicated noise gives identical costs; argmin of
    kernel costs equals a torch argmin; cross-check a 512-candidate slice against the oracle."""
    from ndivplanning_b200 import FusedPlanner
    K, T, nd, E, H = 65536, 8, 2, 128, 1024
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(1, K, T, nd, E, seed=5)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision="fp32")
    plan = pl.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=16)
    cost = plan.cost[0]
    assert torch.isfinite(cost).all()
    perm = torch.randperm(K)
    plan2 = pl.plan(e0.cuda(), goal.cuda(), noise[:, :, perm].cuda(), topk=16)
    assert torch.equal(plan2.cost[0], cost[perm.cuda()])
    assert int(perm[int(plan2.best_idx[0])]) == int(plan.best_idx[0])
    assert int(plan.best_idx[0]) == int(torch.argmin(cost))
    assert set(plan.topk_idx[0].tolist()) == set(torch.sort(cost, stable=True).indices[:16].tolist())
    gen, fwd = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    sl = slice(1000, 1512)
    ref = O.mpc_plan(gen, fwd, e0[0].double(), goal[0].double(), noise[:, 0, sl].double())
    np.testing.assert_allclose(cost[sl].cpu().numpy(), ref["cost"].numpy(), rtol=COST_RTOL)
    full64 = None  # whole-K fp64 oracle is too slow for CI; the winner must beat the slice optimum
    assert float(cost.min()) <= float(ref["cost"].min()) * (1 + 1e-4)


# ---- tensor-core modes of K1 (tcgen05): bulk cost accuracy per mode, measured against the fp64 oracle.
# f16x3 (split activations and weights) must meet the north_star cost tolerance; f16 / f16x2 are
# bulk-estimate modes whose selection is made exact by the fp32 refine pass (test_refined_plan_*).
TC_COST_RTOL = {"f16x3": 1e-4, "f16x2": 6e-4, "f16": 1.2e-3}


@pytest.mark.parametrize("mode", ["f16x3", "f16x2", "f16"])
@pytest.mark.parametrize("G,K,T,H,nd", [(1, 128, 2, 128, 2), (2, 300, 4, 256, 2), (1, 1024, 8, 1024, 2),
                                        (3, 130, 3, 384, 5), (1, 257, 5, 2048, 15)])
def test_tensor_rollout_matches_oracle(cuda, mode, G, K, T, H, nd):
    from ndivplanning_b200 import FusedPlanner
    E = 128
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, nd, E, seed=K + T)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision=mode)
    cost, a0, sT = pl.rollout_score(e0.cuda(), goal.cuda(), noise.cuda(), want_state=True)
    g64, f64 = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    rtol = TC_COST_RTOL[mode]
    for g in range(G):
        ref64 = O.mpc_plan(g64, f64, e0[g].double(), goal[g].double(), noise[:, g].double())
        np.testing.assert_allclose(cost[g].cpu().numpy(), ref64["cost"].numpy(), rtol=rtol)
        scale = float(ref64["a0"].abs().max())
        np.testing.assert_allclose(a0[g].cpu().numpy(), ref64["a0"].numpy(), rtol=0, atol=30 * rtol * scale)
        sscale = float(ref64["state_T"].abs().max())
        np.testing.assert_allclose(sT[g].cpu().numpy(), ref64["state_T"].numpy(), rtol=0, atol=20 * rtol * sscale)


@pytest.mark.parametrize("H", [64, 192, 1000, 1])
def test_any_hidden_width(cuda, H):
    """forward_encoder.py:107-114 has no constraint on the hidden width: the library pads it to a multiple of 128 with
    zero units (exact).  fp32 against the fp32 / fp64 oracle; the default mode (f16+refine) selects the same plan with
    every goal verified or exactly re-rolled."""
    from ndivplanning_b200 import FusedPlanner
    E, nd, G, K, T = 128, 2, 2, 700, 3
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, nd, E, seed=H)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision="fp32")
    plan = pl.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=8, want_state=True)
    dflt = FusedPlanner(gen_p, fwd_p, horizon=T)
    assert dflt.precision == "f16+refine"
    pd = dflt.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=8)
    gen, fwd = _oracle_models(gen_p, fwd_p, nd, E, H)
    g64, f64 = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    for g in range(G):
        ref = O.mpc_plan(gen, fwd, e0[g], goal[g], noise[:, g], topk=8)
        ref64 = O.mpc_plan(g64, f64, e0[g].double(), goal[g].double(), noise[:, g].double())
        np.testing.assert_allclose(plan.cost[g].cpu().numpy(), ref64["cost"].numpy(), rtol=COST_RTOL)
        np.testing.assert_allclose(plan.state_T[g].cpu().numpy(), ref64["state_T"].numpy(), rtol=1e-3, atol=1e-4)
        assert int(plan.best_idx[g]) == int(ref["best_idx"]) == int(pd.best_idx[g])
        assert int(pd.verified[g]) in (1, 2)
    # the single-call module semantics go through the padded copies too
    st, ac = torch.randn(5, E, 1, 1), torch.randn(5, 4)
    with torch.no_grad():
        want = f64(st.double(), ac.double())
    got = pl.forward_model(st.cuda(), ac.cuda()).cpu().double()
    np.testing.assert_allclose(got.numpy(), want.numpy(), rtol=1e-4, atol=1e-5)


def test_two_tile_kernel_variants_are_bit_identical(cuda):
    """the bulk kernel's launch variants — time-split last waves (default for un-pipelined rollouts), unsplit
    (NDP_TC2_NOSPLIT=1, what the pipelined ring launches) and the opt-in 2-CTA cluster with a multicast weight stream
    (NDP_TC2_CLUSTER=1) — must produce the same bits: each runs in its own process (the switches are read once)"""
    import os
    import subprocess
    import sys
    import tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from ndivplanning_b200 import FusedPlanner, synthetic\n"
        "out = []\n"
        "for (G, K, T, H) in ((1, 40000, 3, 256), (2, 300, 4, 256), (1, 513, 2, 128)):\n"
        "    gen, fwd = synthetic.make_models(nd=2, E=128, H=H)\n"
        "    e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=7))\n"
        "    c, a, _ = FusedPlanner(gen, fwd, horizon=T, precision='f16').rollout_score(e0, goal, noise)\n"
        "    out.append((c.cpu(), a.cpu()))\n"
        "torch.save(out, sys.argv[1])\n" % root)
    with tempfile.TemporaryDirectory() as d:
        res = {}
        for name, env in (("default", {}), ("nosplit", {"NDP_TC2_NOSPLIT": "1"}), ("cluster", {"NDP_TC2_CLUSTER": "1"}),
                          ("cluster_nosplit", {"NDP_TC2_CLUSTER": "1", "NDP_TC2_NOSPLIT": "1"})):
            pth = os.path.join(d, name + ".pt")
            subprocess.run([sys.executable, "-c", code, pth], env=dict(os.environ, **env), check=True, timeout=300)
            res[name] = torch.load(pth)
        for name in ("nosplit", "cluster", "cluster_nosplit"):
            for (c, a), (c0, a0) in zip(res[name], res["default"]):
                assert torch.isfinite(c0).all()
                assert torch.equal(c, c0) and torch.equal(a, a0), name


def test_tensor_mode_rejects_unsupported_shapes(cuda):
    from ndivplanning_b200 import FusedPlanner, NdpError
    gen_p, fwd_p = synthetic.make_models(nd=16, E=128, H=128, seed=0)
    with pytest.raises(NdpError):
        FusedPlanner(gen_p, fwd_p, horizon=2, precision="f16x3")     # noise_dim > 15: fp32 mode only
    FusedPlanner(gen_p, fwd_p, horizon=2, precision="fp32")
    gen_p, fwd_p = synthetic.make_models(nd=2, E=256, H=128, seed=0)
    with pytest.raises(NdpError):
        FusedPlanner(gen_p, fwd_p, horizon=2, precision="f16+refine")  # E = 256: fp32 mode only
    assert FusedPlanner(gen_p, fwd_p, horizon=2).precision == "fp32"  # and that is what the default picks
    gen_p, fwd_p = synthetic.make_models(nd=2, E=64, H=128, seed=0)
    with pytest.raises(NdpError):
        FusedPlanner(gen_p, fwd_p, horizon=2, precision="fp32")        # E must be 128 or 256


def test_embedding_256_entry_points(cuda):
    """E = 256 through the other entry points: single-call generator / forward model against the torch modules, the
    pre-allocated ring against plan(), plan() against the fp64 oracle"""
    from ndivplanning_b200 import FusedPlanner
    E, H, nd, G, K, T = 256, 256, 2, 2, 300, 3
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H, seed=3)
    pl = FusedPlanner(gen, fwd, horizon=T)
    z = torch.randn(11, 2 * E + nd)
    s, act = torch.randn(9, E, 1, 1), torch.randn(9, 4)
    with torch.no_grad():
        want_a, want_s = gen(z), fwd(s, act)
    np.testing.assert_allclose(pl.generator(z.cuda()).cpu().numpy(), want_a.numpy(), rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(pl.forward_model(s.cuda(), act.cuda()).cpu().numpy(), want_s.numpy(), rtol=1e-4, atol=1e-5)
    ring = pl.ring(G, K, 4, depth=2)
    for i in range(3):
        e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, nd, E, seed=40 + i))
        q = ring.plan(e0, goal, noise).wait()
        w = pl.plan(e0, goal, noise, topk=4)
        for a, b in ((q.best_idx, w.best_idx), (q.best_action, w.best_action), (q.topk_idx, w.topk_idx), (q.cost, w.cost)):
            assert torch.equal(a, b)
        o = O.mpc_plan(*_oracle_models(gen, fwd, nd, E, H, double=True), e0[0].cpu().double(), goal[0].cpu().double(),
                       noise[:, 0].cpu().double())
        np.testing.assert_allclose(w.cost[0].cpu().numpy(), o["cost"].numpy(), rtol=COST_RTOL)


def _check_selection_bound(cost64, best, topk_list, topk, bound):
    """selection protocol with an explicit cost error bound (relative)"""
    order = np.argsort(cost64, kind="stable")
    gap = (cost64[order[1]] - cost64[order[0]]) / cost64[order[0]] if len(order) > 1 else 1.0
    if gap > 8 * bound:
        assert best == order[0]
    else:
        assert cost64[best] <= cost64[order[0]] * (1 + 8 * bound)
    if topk and len(order) > topk:
        bgap = (cost64[order[topk]] - cost64[order[topk - 1]]) / cost64[order[topk - 1]]
        if bgap > 8 * bound:
            assert set(topk_list) == set(order[:topk].tolist())
        else:
            assert max(cost64[i] for i in topk_list) <= cost64[order[topk - 1]] * (1 + 8 * bound)


@pytest.mark.parametrize("G,K,T,H,topk", [(1, 4096, 8, 1024, 16), (3, 300, 3, 256, 8), (2, 100, 2, 128, 4),
                                          (1, 2048, 4, 512, 0)])
def test_refined_plan_matches_oracle(cuda, G, K, T, H, topk):
    """precision "f16+refine": single-pass bulk estimate of all K, split-operand re-roll of the 128 best,
    plan from the refined costs.  The plan must be the fp64 oracle's plan (up to the fp32-grade bound
    of the refine pass), refined costs must meet the north_star tolerance, and the run must verify."""
    from ndivplanning_b200 import FusedPlanner
    E, nd = 128, 2
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, nd, E, seed=K + T)
    pl = FusedPlanner(gen_p, fwd_p, horizon=T, precision="f16+refine")
    plan = pl.plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=topk)
    exact = FusedPlanner(gen_p, fwd_p, horizon=T, precision="fp32").plan(e0.cuda(), goal.cuda(), noise.cuda(), topk=topk)
    g64, f64 = _oracle_models(gen_p, fwd_p, nd, E, H, double=True)
    assert plan.verified.cpu().tolist() == [1] * G
    for g in range(G):
        ref64 = O.mpc_plan(g64, f64, e0[g].double(), goal[g].double(), noise[:, g].double())
        c64 = ref64["cost"].numpy()
        best = int(plan.best_idx[g])
        tk = plan.topk_idx[g].cpu().tolist() if topk else []
        _check_selection_bound(c64, best, tk, topk, 2e-5)
        # refined entries carry fp32-grade costs / actions; the bulk stays a single-pass estimate
        c = plan.cost[g].cpu().numpy()
        for i in set(tk + [best]):
            assert abs(c[i] - c64[i]) <= COST_RTOL * c64[i]
        np.testing.assert_allclose(c, c64, rtol=1.2e-3)
        np.testing.assert_allclose(plan.best_action[g].cpu().numpy(), ref64["a0"][best].numpy(), rtol=1e-4, atol=1e-5)
        # same plan as the fp32 CUDA path wherever that path is itself unambiguous
        order = np.argsort(c64, kind="stable")
        if (c64[order[1]] - c64[order[0]]) / c64[order[0]] > 8 * 2e-5:
            assert best == int(exact.best_idx[g])


@pytest.mark.parametrize("prec,G,K,T,topk", [("f16+refine", 1, 4096, 4, 16), ("fp32", 2, 300, 3, 4), ("f16x3", 1, 1000, 2, 0),
                                             ("f16+refine", 1, 40000, 4, 16), ("f16+refine", 2, 30000, 3, 8)])
def test_plan_ring_equals_plain_plan(cuda, prec, G, K, T, topk):
    """PlanRing (pre-allocated slots, tails on two side streams) must return exactly what FusedPlanner.plan
    returns, for more plans than the ring has slots (slot reuse) and with results read late.  The two large cases have
    more tile pairs (157, 236) than SMs: FusedPlanner.plan rolls the pairs of the last waves in two pieces on different
    SMs (state images handed over through global memory), the ring does not — every cost must still be the same bits."""
    from ndivplanning_b200 import FusedPlanner
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256)
    pl = FusedPlanner(gen, fwd, horizon=T, precision=prec)
    ring = pl.ring(G, K, topk, depth=3)
    inputs = [synthetic.make_plan_inputs(G, K, T, 2, 128, seed=20 + i) for i in range(7)]
    inputs = [tuple(t.cuda() for t in x) for x in inputs]
    got = []
    for i, (e0, goal, noise) in enumerate(inputs):
        p = ring.plan(e0, goal, noise)
        if i >= 1:                                   # read plan i-1 only after plan i has been issued
            q = got[-1][0].wait()
            got[-1] = tuple(t.clone() if t is not None else None for t in (q.best_idx, q.best_action, q.packed_min,
                                                                           q.topk_idx, q.topk_cost, q.verified, q.cost))
        got.append((p,))
    q = got[-1][0].wait()
    got[-1] = tuple(t.clone() if t is not None else None for t in (q.best_idx, q.best_action, q.packed_min, q.topk_idx,
                                                                   q.topk_cost, q.verified, q.cost))
    for (e0, goal, noise), g in zip(inputs, got):
        w = pl.plan(e0, goal, noise, topk=topk, fallback=False)      # the ring leaves unverified goals flagged
        want = (w.best_idx, w.best_action, w.packed_min, w.topk_idx, w.topk_cost, w.verified, w.cost)
        for a, b in zip(g, want):
            assert (a is None) == (b is None)
            if a is not None:
                assert torch.equal(a, b)
    with pytest.raises(Exception):
        ring.plan(inputs[0][0], inputs[0][1], inputs[0][2][:, :, : K - 1])


@pytest.mark.parametrize("prec,E", [("fp32", 128), ("f16+refine", 128), ("fp32", 256)])
def test_closed_loop_sweep_matches_oracle_loop(cuda, prec, E):
    """N1 driver: G goals swept together on the device == the reference's receding-horizon loop
    (mpc_eval.py:112-176, oracle restatement) run goal by goal with the same noise, incl. the truncated
    horizon near the end (:127)."""
    from ndivplanning_b200.closed_loop import ClosedLoopSweep
    G, K, Th, S, nd, H = 3, 256, 3, 5, 2, 256
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H)
    og, of = _oracle_models(gen, fwd, nd, E, H)
    torch.manual_seed(11)
    start, goal = torch.randn(G, E), torch.randn(G, E)
    fut = torch.randn(S, G, E)
    aref = torch.randn(S, G, 4)
    noises = [torch.rand(min(Th, S - s), G, K, nd) for s in range(S)]
    sw = ClosedLoopSweep(gen, fwd, horizon=Th, rollouts=K, precision=prec)
    res = sw.run(start.cuda(), goal.cuda(), S, state_fut=fut.cuda(), actions_ref=aref.cuda(),
                 noise_fn=lambda s, T: noises[s])
    assert res.candidate_steps == G * K * sum(min(Th, S - s) for s in range(S))
    assert res.verified.cpu().min() >= 1          # 1 = proven by the bound, 2 = full split-operand re-roll
    for g in range(G):
        cur = start[g]
        for s in range(S):
            ref = O.mpc_plan(og, of, cur, goal[g], noises[s][:, g])
            assert int(res.best_idx[s, g]) == ref["best_idx"], (g, s)
            np.testing.assert_allclose(res.actions[s, g].cpu().numpy(), ref["best_action"].numpy(), rtol=2e-3, atol=1e-5)
            with torch.no_grad():
                cur = of(cur.view(1, E, 1, 1), ref["best_action"][None]).view(E)
            np.testing.assert_allclose(res.states[s, g].cpu().numpy(), cur.numpy(), rtol=2e-3, atol=2e-4)
            np.testing.assert_allclose(float(res.image_err[s, g]), float(((cur - fut[s, g]) ** 2).mean()), rtol=2e-3)
    want_ae = ((aref - res.actions.cpu()) ** 2).mean(dim=(0, 2))
    np.testing.assert_allclose(res.action_err.cpu().numpy(), want_ae.numpy(), rtol=1e-5)
    # device sampler path: runs without host noise and is reproducible under a seeded device generator
    g1 = torch.Generator(device="cuda").manual_seed(5)
    g2 = torch.Generator(device="cuda").manual_seed(5)
    r1 = sw.run(start.cuda(), goal.cuda(), S, generator=g1)
    r2 = sw.run(start.cuda(), goal.cuda(), S, generator=g2)
    assert torch.equal(r1.best_idx, r2.best_idx) and torch.equal(r1.actions, r2.actions)


@pytest.mark.parametrize("prec", ["f16+refine", "fp32"])
def test_closed_loop_sweep_as_one_cuda_graph(cuda, prec):
    """the whole receding-horizon loop captured as ONE CUDA graph: replays with refilled static inputs
    equal the eager sweep bit for bit; with the in-graph device sampler every replay draws fresh noise"""
    from ndivplanning_b200.closed_loop import ClosedLoopSweep
    G, K, Th, S, nd, E, H = 4, 512, 3, 4, 2, 128, 256
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=H)
    torch.manual_seed(12)
    start, goal = torch.randn(G, E).cuda(), torch.randn(G, E).cuda()
    fut, aref = torch.randn(S, G, E).cuda(), torch.randn(S, G, 4).cuda()
    noises = [torch.rand(min(Th, S - s), G, K, nd).cuda() for s in range(S)]
    sw = ClosedLoopSweep(gen, fwd, horizon=Th, rollouts=K, precision=prec)
    graph, res = sw.capture(start, goal, S, state_fut=fut, actions_ref=aref, noise_fn=lambda s, T: noises[s])
    for trial in range(2):                                   # second trial: new goals and noise in the same buffers
        if trial:
            goal.copy_(torch.randn(G, E))
            for n_ in noises:
                n_.copy_(torch.rand(n_.shape))
        graph.replay()
        torch.cuda.synchronize()
        got = [t.clone() for t in (res.best_idx, res.actions, res.states, res.image_err, res.action_err, res.verified)]
        ref = sw.run(start, goal, S, state_fut=fut, actions_ref=aref, noise_fn=lambda s, T: noises[s])
        want = (ref.best_idx, ref.actions, ref.states, ref.image_err, ref.action_err, ref.verified)
        for a, b in zip(got, want):
            assert torch.equal(a, b)
    g2, r2 = sw.capture(start, goal, S)                      # device sampler inside the graph
    g2.replay(); torch.cuda.synchronize()
    first = r2.actions.clone()
    g2.replay(); torch.cuda.synchronize()
    assert not torch.equal(first, r2.actions) and bool(torch.isfinite(r2.actions).all())


def test_refine_fallback_rerolls_unverified_goals(cuda):
    """Goals whose refined selection cannot be proven exact (costs too close: short horizon, weak separation)
    are re-rolled over all K with split operands on the device: verified == 2 and the goal's costs, argmin and
    top-k are exactly those of the f16x3 planner; verified == 1 goals keep the refined plan."""
    from ndivplanning_b200 import FusedPlanner
    G, K, T, topk = 8, 512, 1, 8
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256, sep=1.0)
    e0, goal, noise = synthetic.make_plan_inputs(G, K, T, 2, 128, seed=4)
    e0, goal, noise = e0.cuda(), goal.cuda(), noise.cuda()
    plan = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine").plan(e0, goal, noise, topk=topk)
    full = FusedPlanner(gen, fwd, horizon=T, precision="f16x3").plan(e0, goal, noise, topk=topk)
    ver = plan.verified.cpu().tolist()
    assert set(ver) <= {1, 2} and 2 in ver, ver
    for g in range(G):
        if ver[g] == 2:
            assert torch.equal(plan.cost[g], full.cost[g])
            assert int(plan.best_idx[g]) == int(full.best_idx[g])
            assert torch.equal(plan.topk_idx[g], full.topk_idx[g]) and torch.equal(plan.topk_cost[g], full.topk_cost[g])
            assert torch.equal(plan.best_action[g], full.best_action[g])
        else:
            assert set(plan.topk_idx[g].tolist()) == set(full.topk_idx[g].tolist())


def test_ring_fallback_consumer_initiated(cuda):
    """PlanRing leaves unverified goals flagged (verified == 0); ring.fallback(plan) turns them into the exact plan"""
    from ndivplanning_b200 import FusedPlanner
    G, K, T, topk = 8, 512, 1, 8
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256, sep=1.0)
    e0, goal, noise = (t.cuda() for t in synthetic.make_plan_inputs(G, K, T, 2, 128, seed=4))
    pl = FusedPlanner(gen, fwd, horizon=T, precision="f16+refine")
    ring = pl.ring(G, K, topk)
    p = ring.plan(e0, goal, noise).wait()
    torch.cuda.synchronize()
    assert 0 in p.verified.cpu().tolist()
    ring.fallback(p).wait()
    want = pl.plan(e0, goal, noise, topk=topk)
    torch.cuda.synchronize()
    assert torch.equal(p.verified, want.verified) and 2 in p.verified.cpu().tolist()
    assert torch.equal(p.best_idx, want.best_idx) and torch.equal(p.best_action, want.best_action)
    assert torch.equal(p.topk_idx, want.topk_idx) and torch.equal(p.cost, want.cost)


def test_device_sampler_matches_philox_oracle(cuda):
    """ndp_sample_uniform (row A on the device) is bit-exact against the numpy Philox4x32-10 oracle, which is pinned
    by the Random123 known-answer vectors (tests/test_oracle_golden.py); and it drives a plan"""
    from ndivplanning_b200 import FusedPlanner, sample_noise_device
    for seed, offset, shape in ((0, 0, (1, 1, 5, 3)), (1234567890123456789, 2 ** 40 + 17, (2, 3, 257, 2)),
                                (2 ** 64 - 1, 2 ** 63, (8, 1, 4096, 2))):
        got = sample_noise_device(*shape, seed=seed, offset=offset, device=cuda).cpu().numpy().reshape(-1)
        want = O.sample_uniform(seed, offset, got.size)
        assert np.array_equal(got, want), (seed, offset)
        assert got.min() >= 0.0 and got.max() < 1.0
    a = sample_noise_device(4, 1, 1024, 2, seed=5, offset=0, device=cuda)
    b = sample_noise_device(4, 1, 1024, 2, seed=5, offset=1, device=cuda)
    assert not torch.equal(a, b)
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256)
    e0, goal, _ = synthetic.make_plan_inputs(1, 1024, 4, 2, 128, seed=2)
    pl = FusedPlanner(gen, fwd, horizon=4, precision="fp32")
    p1 = pl.plan(e0.cuda(), goal.cuda(), a)
    og, of = _oracle_models(gen, fwd, 2, 128, 256)
    ref = O.mpc_plan(og, of, e0[0], goal[0], a.cpu()[:, 0])
    assert int(p1.best_idx[0]) == ref["best_idx"]
