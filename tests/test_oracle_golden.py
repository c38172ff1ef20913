"""CPU: the oracle (torch-CPU restatement) against the golden vectors produced by the
unmodified reference (oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import ndiv_oracle as O
from ndivplanning_b200 import synthetic

NDIV_CASES = ["ref_shape", "k2", "k3", "k64", "k256", "k100_ragged"]


@pytest.mark.parametrize("name", NDIV_CASES)
def test_ndiv_oracle_matches_reference(golden, name):
    g = golden["ndiv"]
    x, z = torch.from_numpy(g[f"{name}_x"]), torch.from_numpy(g[f"{name}_z"])
    loss, grad = O.ndiv_fwd_bwd(x, z)
    np.testing.assert_allclose(loss.numpy(), g[f"{name}_loss"], rtol=1e-6)
    np.testing.assert_allclose(grad.numpy(), g[f"{name}_grad"], rtol=1e-5, atol=1e-8)
    # closed-form gradient (what the kernel implements) vs the reference's autograd, fp64
    ga = O.ndiv_grad_analytic(x.double(), z.double())
    np.testing.assert_allclose(ga.numpy(), g[f"{name}_grad64"], rtol=1e-9, atol=1e-12)
    if f"{name}_pairwise" in g:
        np.testing.assert_allclose(O.compute_pairwise(x).numpy(), g[f"{name}_pairwise"], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(O.compute_pair_distance(x).numpy(), g[f"{name}_pairdist"], rtol=1e-6, atol=1e-8)


def test_ndiv_k1_is_nan():
    x, z = torch.randn(2, 1, 4), torch.rand(2, 1, 2)
    assert torch.isnan(O.compute_pairwise_divergence(x, z))


@pytest.mark.parametrize("nd", [2, 16])
def test_generator_oracle_matches_reference(golden, nd):
    g = golden["gen"]
    torch.manual_seed(0)
    gen = O.GeneratorMLP(nd).eval()
    with torch.no_grad():
        a = gen(torch.from_numpy(g[f"nd{nd}_z"]))
    np.testing.assert_allclose(a.numpy(), g[f"nd{nd}_a"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["small", "cfg1"])
def test_mpc_oracle_matches_reference(golden, name):
    g = golden["mpc"]
    K, Th, seq, n_traj, E, H, nd, seed = [int(v) for v in g[f"{name}_params"]]
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
    gen, fwd = O.GeneratorMLP(nd).eval(), O.LatentForward(E, H).eval()
    gen.load_state_dict(gen_p.state_dict()); fwd.load_state_dict(fwd_p.state_dict())
    images, actions = torch.from_numpy(g[f"{name}_images"]), torch.from_numpy(g[f"{name}_actions"])
    a, i, best = O.closed_loop_eval(gen, fwd, images, actions, K, Th, nd, seed=seed)
    np.testing.assert_allclose([a, i], g[f"{name}_tuple"], rtol=1e-6)
    assert np.array_equal(np.array(best), g[f"{name}_best"])
    e0, goal, noise = synthetic.make_plan_inputs(1, K, Th, nd, E, seed=21)
    plan = O.mpc_plan(gen, fwd, e0[0], goal[0], noise[:, 0], topk=8)
    np.testing.assert_allclose(plan["cost"].numpy(), g[f"{name}_plan_cost"], rtol=1e-5)
    assert plan["best_idx"] == int(g[f"{name}_plan_best"])
    assert set(plan["topk_idx"].tolist()) == set(g[f"{name}_plan_topk"].tolist())


def test_select_first_min_rules():
    inf, nan = float("inf"), float("nan")
    assert O.select_first_min(torch.tensor([3.0, 1.0, 1.0, 2.0])) == 1        # first minimal index
    assert O.select_first_min(torch.tensor([nan, 5.0, nan, 4.0])) == 3        # NaN never wins
    assert O.select_first_min(torch.tensor([1e10, 2e10, inf])) == 0           # nothing < 1e10 -> 0
    assert O.select_first_min(torch.tensor([nan, nan])) == 0
    assert O.select_topk(torch.tensor([2.0, 1.0, nan, 1.0, 0.5]), 3).tolist() == [4, 1, 3]
    assert O.pack_key(0.5, 7) < O.pack_key(0.75, 3) < O.pack_key(float("nan"), 0)
    assert O.pack_key(-1.0, 9) < O.pack_key(0.0, 0)


GAN_CASES = ["ref_shape", "s16", "nd4", "nodiv"]


def _flat(m, grad=False):
    return torch.cat([(p.grad if grad else p.detach()).reshape(-1) for p in m.parameters()])


@pytest.mark.parametrize("name", GAN_CASES)
def test_gan_step_oracle_matches_reference(golden, name):
    """restated train_gan.py:126-203 loop body (oracle containers) vs the run with the reference's own modules"""
    g = golden["gan"]
    Fb, S, nd, dsteps, nsteps, seed = [int(v) for v in g[f"{name}_cfg"]]
    torch.manual_seed(0)
    gen, dis = O.GeneratorMLP(nd).train(), O.ActionDiscriminator().train()
    G_opt, D_opt = O.gan_optimizers(gen, dis, lr=2e-4)
    losses = []
    for it in range(nsteps):
        codes, actions, noise = synthetic.make_gan_batch(Fb, S, nd, seed=seed + 100 * it)
        probe = {}
        dl, gl, pl, a_hat = O.gan_train_step(gen, dis, G_opt, D_opt, codes, actions, noise, float(g[f"{name}_divf"]),
                                             dsteps, probe=probe)
        losses.append([float(dl), float(gl), float(pl)])
        if it == 0:
            np.testing.assert_allclose(a_hat.numpy(), g[f"{name}_a_hat0"], rtol=1e-5, atol=1e-7)
            if name == "ref_shape":
                np.testing.assert_allclose(_flat(gen, True).numpy(), g[f"{name}_ggrad0"], rtol=1e-4, atol=1e-8)
                np.testing.assert_allclose(probe["d_grad"].numpy(), g[f"{name}_dgrad0"], rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(np.array(losses), g[f"{name}_losses"], rtol=1e-5)
    st = 1 if name == "ref_shape" else 4        # GAN_STRIDE64 of oracle/make_golden.py
    np.testing.assert_allclose(_flat(gen).numpy()[::st], g[f"{name}_gparams"], rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(_flat(dis).numpy()[::st], g[f"{name}_dparams"], rtol=1e-4, atol=2e-6)


def test_philox_known_answers():
    """Philox4x32-10 known-answer vectors of the Random123 distribution (kat_vectors): the oracle of the device sampler"""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = O.philox4x32_10([ctr], key)[0]
        assert tuple(int(v) for v in got) == want, (ctr, [hex(int(v)) for v in got])
    u = O.sample_uniform(7, 3, 10001)
    assert u.shape == (10001,) and u.min() >= 0.0 and u.max() < 1.0 and abs(float(u.mean()) - 0.5) < 0.02


@pytest.mark.parametrize("name", ["b1", "b2", "b5"])
def test_encoder_oracle_matches_reference(golden, name):
    """oracle.encoder_forward vs the reference Encoder (models/image_autoencoder.py:35-49, eval mode) run by
    oracle/make_golden.py::golden_enc on the same seeded frames and parameters."""
    g = golden["enc"]
    B, seed, pseed = (int(v) for v in g[f"{name}_cfg"])
    enc = synthetic.make_image_encoder(seed=pseed)
    psum = sum(float(v.double().sum()) for v in enc.state_dict().values())
    assert abs(psum - float(g["param_sum"])) < 1e-9 * abs(float(g["param_sum"])) + 1e-9     # same RNG stream as the fixture
    frames = synthetic.make_frames(B, seed=seed)
    assert abs(float(frames.double().sum()) - float(g[f"{name}_frame_sum"])) < 1e-6
    with torch.no_grad():
        y64 = O.encoder_forward(enc.state_dict(), frames.double()).view(B, 128).numpy()
        y32 = O.encoder_forward(enc.state_dict(), frames).view(B, 128).numpy()
        ymod = enc(frames).view(B, 128).numpy()          # the container's reference-style forward
    ref64, ref32 = g[f"{name}_codes64"], g[f"{name}_codes"]
    assert np.abs(y64 - ref64).max() < 1e-12
    assert np.abs(y32 - ref64).max() < 4 * np.abs(ref32 - ref64).max() + 1e-6
    assert np.abs(ymod - ref32).max() < 2e-6


def test_encoder_fold_matches_unfolded():
    """BatchNorm folding (ndivplanning_b200/encoder.py::fold_state_dict) is an identity on the eval-mode forward."""
    from ndivplanning_b200.encoder import fold_state_dict
    enc = synthetic.make_image_encoder(seed=11)
    frames = synthetic.make_frames(2, seed=5).double()
    with torch.no_grad():
        want = O.encoder_forward(enc.state_dict(), frames)
        x = frames
        for i, (W, b) in enumerate(fold_state_dict({k: v.double() for k, v in enc.state_dict().items()})):
            W = W.double().permute(0, 3, 1, 2)        # back to [Cout,Cin,KH,KW]
            x = torch.nn.functional.conv2d(x, W, b.double(), stride=2 if i < 5 else 1, padding=1 if i < 5 else 0)
            if i < 5:
                x = torch.relu(x)
    assert float((x - want).abs().max()) < 2e-6       # folded weights are rounded to fp32 once


def test_chunked_ndiv_equals_autograd():
    """the row-blocked fp64 evaluation used as the arbiter at BASELINE cfg3 size equals diversity.py under autograd
    (restated in oracle.compute_pairwise_divergence, itself pinned on the reference above), incl. d loss / d codes"""
    torch.manual_seed(3)
    x = torch.randn(1, 300, 12, dtype=torch.float64, requires_grad=True)
    z = torch.rand(1, 300, 5, dtype=torch.float64, requires_grad=True)
    loss = O.compute_pairwise_divergence(x, z)
    loss.backward()
    l2, gx, sx, sz, gz = O.ndiv_fwd_bwd_chunked(x.detach()[0], z.detach()[0], block=64, want_grad_z=True)
    np.testing.assert_allclose(l2.item(), loss.item(), rtol=1e-12)
    np.testing.assert_allclose(gx.numpy(), x.grad[0].numpy(), rtol=1e-9, atol=1e-14)
    np.testing.assert_allclose(gz.numpy(), z.grad[0].numpy(), rtol=1e-9, atol=1e-14)
    np.testing.assert_allclose(sx.numpy(), O.compute_pairwise(x.detach())[0].sum(1).numpy(), rtol=1e-12)
