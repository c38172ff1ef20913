"""GPU parity of the device-resident GAN training step (csrc/k4_gan_train.cu through
ndp_gan_train_step) against tests/golden/gan.npz — train_gan.py:126-203 run with the reference's own
Decoder / Discriminator / diversity modules and torch.optim.Adam — and against the oracle restatement.

Protocol.  Adam's first update is sign-like (m/sqrt(v) = +-1): a parameter whose gradient is within round-off of
zero moves by +-lr depending on the last bit, in ANY fp32 implementation (the reference's own fp32 and fp64 runs
differ by 4e-4 = 2*lr in a few parameters after one step, golden).  With the ndiv term on, those jumps are then
amplified chaotically (the S samples of a code sit ~1e-3 apart, the normalised distances and the hinge mask react to
1e-4 shifts): after two steps the reference's fp32 and fp64 generator gradients differ by 6 % of their scale.  So:
  * the FIRST step — forward, all three losses, both gradients — is compared tightly (arbitrated by the fp64 run);
  * the optimiser recurrence over several steps is compared tightly on the 'nodiv' case (pairwise_div_factor = 0),
    where the reference's fp32 run stays within 5e-7 of its fp64 run;
  * multi-step runs WITH the ndiv term must stay within a multiple of the reference's own fp32-vs-fp64 drift."""
import numpy as np
import pytest
import torch

from ndivplanning_b200 import GanTrainer, synthetic
from ndivplanning_b200.gan_train import D_LAYERS, G_LAYERS
from oracle import ndiv_oracle as O

pytestmark = pytest.mark.gpu
S64 = 4     # stride of the fp64 parameter samples in the fixture (oracle/make_golden.py GAN_STRIDE64)


def _flat_sd(sd, layers):
    return torch.cat([sd[f"{n}.{k}"].reshape(-1) for n in layers for k in ("weight", "bias")]).cpu().numpy()


def _arbiter(mine, ref32, ref64, what, floor=1e-7, k=4):
    e_m, e_r = np.abs(mine.astype(np.float64) - ref64), np.abs(ref32.astype(np.float64) - ref64)
    assert e_m.max() <= max(k * e_r.max(), 50 * floor), (what, "max", e_m.max(), e_r.max())
    rms_m, rms_r = np.sqrt((e_m ** 2).mean()), np.sqrt((e_r ** 2).mean())
    assert rms_m <= k * rms_r + floor, (what, "rms", rms_m, rms_r)
    thr = 100 * floor
    assert (e_m > thr).mean() <= k * (e_r > thr).mean() + 1e-3, (what, "outliers", (e_m > thr).mean(), (e_r > thr).mean())
    assert np.median(e_m) <= k * np.median(e_r) + floor, (what, "median", np.median(e_m), np.median(e_r))


def _run(cuda, name, golden, graph=False):
    g = golden["gan"]
    Fb, S, nd, dsteps, nsteps, seed = [int(v) for v in g[f"{name}_cfg"]]
    gen, dis = synthetic.make_gan_models(nd, seed=0)
    tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=Fb, learning_rate=2e-4, betas=(0.5, 0.999),
                    pairwise_div_factor=float(g[f"{name}_divf"]), discrim_steps_per_gen=dsteps, device=cuda)
    losses, extra = [], {}
    static = [t.to(cuda) for t in synthetic.make_gan_batch(Fb, S, nd, seed=seed)]
    cg = tr.capture(*static) if graph else None
    if graph:           # capture does not execute; reset nothing, the first replay is step 1
        pass
    for it in range(nsteps):
        batch = [t.to(cuda) for t in synthetic.make_gan_batch(Fb, S, nd, seed=seed + 100 * it)]
        if graph:
            for dst, src in zip(static, batch):
                dst.copy_(src)
            cg.replay()
            l = tr.losses
        else:
            l = tr.step(*batch)
        losses.append(l.cpu().numpy().astype(np.float64))
        if it == 0:
            extra["a_hat0"] = tr.last_actions(Fb).cpu().numpy()
            extra["ggrad0"], extra["dgrad0"] = [t.cpu().numpy() for t in tr.grads()]
    gsd, dsd = tr.state_dicts()
    return np.array(losses), _flat_sd(gsd, G_LAYERS), _flat_sd(dsd, D_LAYERS), extra


@pytest.mark.parametrize("name", ["ref_shape", "s16", "nd4", "nodiv"])
def test_gan_step_matches_reference_run(cuda, golden, name):
    g = golden["gan"]
    losses, gp, dp, extra = _run(cuda, name, golden)
    st = 1 if name == "ref_shape" else S64
    # ---- first step: plain fp32 arithmetic on identical parameters
    np.testing.assert_allclose(extra["a_hat0"], g[f"{name}_a_hat0"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(losses[0, 0], g[f"{name}_losses"][0, 0], rtol=1e-5)
    np.testing.assert_allclose(losses[0, 1], g[f"{name}_losses"][0, 1], rtol=2e-5)
    np.testing.assert_allclose(losses[0, 2], g[f"{name}_losses"][0, 2], rtol=1e-4)
    if name == "ref_shape":
        for key in ("ggrad0", "dgrad0"):
            r32, r64 = g[f"{name}_{key}"], g[f"{name}_{key}64"]
            scale = np.abs(r64).max()
            _arbiter(extra[key] / scale, r32 / scale, r64 / scale, key, floor=1e-7)
    # ---- all steps
    l32, l64 = g[f"{name}_losses"], g[f"{name}_losses64"]
    smooth = name == "nodiv"
    k = 4 if smooth else 12
    ref_err = np.abs(l32 - l64).max(axis=0, keepdims=True)
    assert np.all(np.abs(losses - l64) <= k * ref_err + (2e-6 if smooth else 1e-4) * np.abs(l64)), (losses, l32, l64)
    _arbiter(gp[::S64], g[f"{name}_gparams"][::S64 // st], g[f"{name}_gparams64"], "generator parameters", k=k)
    _arbiter(dp[::S64], g[f"{name}_dparams"][::S64 // st], g[f"{name}_dparams64"], "discriminator parameters", k=k)


def test_gan_step_graph_replay_is_bit_identical(cuda, golden):
    """the step is deterministic (no atomics) and stream-ordered: CUDA-graph replays == eager steps, bit for bit"""
    a = _run(cuda, "ref_shape", golden, graph=False)
    b = _run(cuda, "ref_shape", golden, graph=True)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_gan_step_large_batch_vs_oracle(cuda):
    """F=2048 codes x S=16 samples (32,768 generator rows): one step against the oracle restatement on the CPU"""
    Fb, S, nd = 2048, 16, 2
    gen, dis = synthetic.make_gan_models(nd, seed=1)
    codes, actions, noise = synthetic.make_gan_batch(Fb, S, nd, seed=9)
    tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=Fb, device=cuda)
    l = tr.step(codes.to(cuda), actions.to(cuda), noise.to(cuda)).cpu().numpy()
    gg, dg = [t.cpu().numpy() for t in tr.grads()]
    res = {}
    for dt in (torch.float32, torch.float64):
        og, od = O.GeneratorMLP(nd).to(dt), O.ActionDiscriminator().to(dt)
        og.load_state_dict({k: v.to(dt) for k, v in gen.state_dict().items()})
        od.load_state_dict({k: v.to(dt) for k, v in dis.state_dict().items()})
        G_opt, D_opt = O.gan_optimizers(og, od)
        # discriminator gradient of the D step: read it before the generator backward overwrites .grad
        dl, gl, pl, _ = O.gan_train_step(og, od, G_opt, D_opt, codes.to(dt), actions.to(dt), noise.to(dt))
        res[dt] = (np.array([float(dl), float(gl), float(pl)]),
                   torch.cat([p.grad.reshape(-1) for p in og.parameters()]).numpy())
    l32, l64 = res[torch.float32][0], res[torch.float64][0]
    assert np.all(np.abs(l - l64) <= 4 * np.abs(l32 - l64) + 1e-4 * np.abs(l64)), (l, l32, l64)
    g32, g64 = res[torch.float32][1], res[torch.float64][1]
    scale = np.abs(g64).max()
    # ill-conditioned through the ndiv term (see the tensor-path test): the tcgen05 products carry 2^-21 instead of
    # 2^-24 per term, so allow a wider multiple of the reference's own fp32-vs-fp64 error here
    _arbiter(gg / scale, g32 / scale, g64 / scale, "generator gradient (large)", floor=1e-7, k=10)


def test_gan_trainer_refuses_cpu_and_bad_shapes(cuda):
    gen, dis = synthetic.make_gan_models(2, seed=0)
    tr = GanTrainer(gen, dis, num_sample=4, noise_dim=2, max_codes=8, device=cuda)
    c, a, n = synthetic.make_gan_batch(8, 4, 2)
    with pytest.raises(Exception):
        tr.step(c, a, n)                                   # CPU tensors: no fallback
    with pytest.raises(Exception):
        tr.step(c.to(cuda)[:, :100], a.to(cuda), n.to(cuda))
    c16, a16, n16 = [t.to(cuda) for t in synthetic.make_gan_batch(16, 4, 2)]
    with pytest.raises(Exception):
        tr.step(c16, a16, n16)                             # F > max_codes


@pytest.mark.parametrize("divf", [0.0, 0.1])
def test_gan_step_tensor_path_matches_exact_path(cuda, monkeypatch, divf):
    """wide layers at large batch run on tcgen05 with split-fp16 operands (22-bit products, power-of-two operand
    scales); NDP_GAN_EXACT=1 keeps the same step on the fp32 FFMA kernels.  Without the ndiv term every quantity of
    the first step must agree to fp32 grade.  With it the generator gradient is ill-conditioned (the S samples of a
    code sit ~1e-3 apart: the reference's own fp32 run is 1e-4 rms / 3e-3 max of the gradient scale away from its
    fp64 run, see test_gan_step_large_batch_vs_oracle), so only the same order of agreement is asked for."""
    Fb, S, nd = 3000, 8, 2          # 24,000 generated rows (ragged against the 128-row tiles), 3000 real rows
    gen, dis = synthetic.make_gan_models(nd, seed=2)
    batch = [t.to(cuda) for t in synthetic.make_gan_batch(Fb, S, nd, seed=11)]
    out = {}
    for mode in ("tc", "exact"):
        if mode == "exact":
            monkeypatch.setenv("NDP_GAN_EXACT", "1")
        else:
            monkeypatch.delenv("NDP_GAN_EXACT", raising=False)
        tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=Fb, pairwise_div_factor=divf, device=cuda)
        l = tr.step(*batch).cpu().numpy().astype(np.float64)
        gg, dg = [t.cpu().numpy().astype(np.float64) for t in tr.grads()]
        out[mode] = (l, tr.last_actions(Fb).cpu().numpy().astype(np.float64), gg, dg)
    np.testing.assert_allclose(out["tc"][0][:2], out["exact"][0][:2], rtol=2e-6)
    np.testing.assert_allclose(out["tc"][0][2], out["exact"][0][2], rtol=1e-4)        # hinge of near-coincident samples
    np.testing.assert_allclose(out["tc"][1], out["exact"][1], rtol=1e-5, atol=2e-7)
    for k, name in ((2, "generator"), (3, "discriminator")):
        a, b = out["tc"][k], out["exact"][k]
        err = np.abs(a - b) / np.abs(b).max()
        rms = np.sqrt((err ** 2).mean())
        if name == "generator" and divf > 0:
            assert rms < 1e-3 and np.median(err) < 1e-4, (name, err.max(), rms)
        else:
            # 22-bit products through a chain of up to eight layer products (the discriminator's chain is three:
            # median 2e-8); isolated larger entries are (leaky-)ReLU derivatives of units whose pre-activation is
            # within round-off of 0 and lands on the other side
            assert rms < 1e-5 and np.median(err) < 2e-6 and err.max() < 2e-3, (name, err.max(), rms, np.median(err))


def test_gan_step_edge_shapes_and_lr(cuda):
    """one code, two samples; zero discriminator steps; learning-rate change between steps — against the oracle"""
    for Fb, S, nd, dsteps in ((1, 2, 2, 1), (3, 4, 15, 0), (5, 32, 3, 1)):
        gen, dis = synthetic.make_gan_models(nd, seed=4)
        c, a, z = synthetic.make_gan_batch(Fb, S, nd, seed=6)
        tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=max(Fb, 8), discrim_steps_per_gen=dsteps, device=cuda)
        og, od = O.GeneratorMLP(nd), O.ActionDiscriminator()
        og.load_state_dict(gen.state_dict()); od.load_state_dict(dis.state_dict())
        G_opt, D_opt = O.gan_optimizers(og, od)
        for it, lr in enumerate((2e-4, 5e-5)):
            tr.set_lr(lr)
            for opt in (G_opt, D_opt):
                for grp in opt.param_groups:
                    grp["lr"] = lr
            l = tr.step(c.to(cuda), a.to(cuda), z.to(cuda)).cpu().numpy()
            dl, gl, pl, _ = O.gan_train_step(og, od, G_opt, D_opt, c, a, z, 0.1, dsteps)
            if dsteps:
                np.testing.assert_allclose(l[0], float(dl), rtol=1e-5)
            np.testing.assert_allclose(l[1], float(gl), rtol=1e-4 if it else 2e-5)
            np.testing.assert_allclose(l[2], float(pl), rtol=2e-3 if it else 1e-4)
        gsd, dsd = tr.state_dicts()
        gp = _flat_sd(gsd, G_LAYERS)
        rp = torch.cat([p.detach().reshape(-1) for p in og.parameters()]).numpy()
        # lr-sized sign jumps of parameters with round-off-sized gradients are allowed (see the module docstring)
        assert np.median(np.abs(gp - rp)) < 1e-7 and np.abs(gp - rp).max() <= 2.2 * (2e-4 + 5e-5), (Fb, S, np.abs(gp - rp).max())


def test_gan_single_sample_groups_are_nan_like_the_reference(cuda):
    """num_sample = 1: the normalised distance of a one-point group is 0/0 (diversity.py:17-18) -> NaN loss"""
    gen, dis = synthetic.make_gan_models(2, seed=0)
    c, a, z = synthetic.make_gan_batch(4, 1, 2, seed=6)
    tr = GanTrainer(gen, dis, num_sample=1, noise_dim=2, max_codes=4, device=cuda)
    l = tr.step(c.to(cuda), a.to(cuda), z.to(cuda)).cpu().numpy()
    assert np.isfinite(l[0]) and np.isfinite(l[1]) and np.isnan(l[2])


class _Traj(torch.utils.data.Dataset):
    """duck-typed PushDataset (utils/trajectory_loader.py:38-72) in embedding space: images are [seq,E,1,1] codes"""

    def __init__(self, n, seq, E=128, seed=0):
        g = torch.Generator().manual_seed(seed)
        self.images = torch.randn(n, seq, E, 1, 1, generator=g)
        self.actions = torch.randn(n, seq, 4, generator=g)
        self.seq_length = seq

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], torch.zeros(self.seq_length, 25), self.actions[i], torch.zeros(3)


def test_train_gan_epoch_matches_oracle_epoch(cuda):
    """the epoch loop around the step (train_gan.py:114-209: unpack trajectories, encode, sample, step, average the
    losses) against the oracle's CPU epoch with the same batches and noise; pairwise_div_factor = 0 keeps the
    comparison out of the chaotic regime (see the module docstring)"""
    from ndivplanning_b200 import train_gan_epoch
    seq, bs, S, nd = 8, 8, 6, 2
    ds = _Traj(24, seq)
    loader = torch.utils.data.DataLoader(ds, batch_size=bs, shuffle=False)
    enc = torch.nn.Identity()
    F = bs * (seq - 1)
    torch.manual_seed(3)
    noises = [torch.rand(F, S, nd) for _ in range(len(loader))]
    gen, dis = synthetic.make_gan_models(nd, seed=0)
    tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=F, pairwise_div_factor=0.0, device=cuda)
    it = iter(noises)
    got = train_gan_epoch(tr, loader, enc, seq, noise_fn=lambda F_: next(it))
    og, od = O.GeneratorMLP(nd), O.ActionDiscriminator()
    og.load_state_dict(gen.state_dict()); od.load_state_dict(dis.state_dict())
    G_opt, D_opt = O.gan_optimizers(og, od)
    want = O.gan_train_epoch(og, od, G_opt, D_opt, loader, enc, seq, S, nd, div_factor=0.0, noises=noises)
    np.testing.assert_allclose(got[:2], want[:2], rtol=2e-5)
    np.testing.assert_allclose(got[2], want[2], rtol=2e-3)
    gsd, _ = tr.state_dicts()
    rp = torch.cat([p.detach().reshape(-1) for p in og.parameters()]).numpy()
    assert np.median(np.abs(_flat_sd(gsd, G_LAYERS) - rp)) < 1e-7
    # device sampler path runs too
    d2 = train_gan_epoch(tr, loader, enc, seq)
    assert all(np.isfinite(v) for v in d2)
