"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/*.npz by running the UNMODIFIED
reference (imported from /root/reference through oracle/ref_stubs.py) in the build
container.  The reference cannot travel to the GPU box; the vectors do.

    python oracle/make_golden.py            # rewrites tests/golden/

What is pinned:
  ndiv_*.npz   diversity.compute_pairwise_divergence (+autograd grad, compute_pairwise,
               compute_pair_distance) on seeded inputs, incl. the reference call shape
               N=56,k=6,Dx=4,Dz=2 (train_gan.py:193-196), duplicates, k=2/k=3 cases.
  gen_*.npz    models.gan.Decoder forward on seeded inputs + proof that the product's
               GeneratorMLP gets bit-identical default-init weights under the same seed.
  gan.npz      train_gan.py:126-203 loop body run with the reference's own models.gan.Decoder / Discriminator,
               diversity.compute_pairwise_divergence and torch.optim.Adam (:98-104): per-step losses, first-step
               action_hat and generator gradient, parameters after the last step; fp32 and fp64.
  mpc_*.npz    mpc_eval.fetch_push_control_evaluation (unmodified, embedding mode:
               nn.Identity encoder) return tuple for a fake in-memory dataset, plus
               per-plan costs/argmin from the reference's modules via the restated loop.
"""
import contextlib
import io
import logging
import os
import sys

import numpy as np
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ndiv_oracle as O  # noqa: E402
from oracle.ref_stubs import DotMap, import_reference  # noqa: E402
from ndivplanning_b200 import synthetic  # product-side input spec only (no compute)  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def golden_ndiv(div):
    cases = {
        "ref_shape": (56, 6, 4, 2, 10),     # train_gan.py:168-169,193-196
        "k2": (3, 2, 4, 2, 11),
        "k3": (2, 3, 5, 3, 12),
        "k64": (2, 64, 32, 16, 13),
        "k256": (1, 256, 32, 16, 14),       # BASELINE config 1 ndiv batch
        "k100_ragged": (3, 100, 7, 3, 15),
    }
    out = {}
    for name, (N, k, Dx, Dz, seed) in cases.items():
        x, z = synthetic.make_ndiv_inputs(N, k, Dx, Dz, seed=seed)
        if name == "k64":  # duplicates: zero distances off the diagonal (finite grads)
            x[:, 5] = x[:, 9]
            z[:, 7] = z[:, 3]
        xr = x.clone().requires_grad_(True)
        loss = div.compute_pairwise_divergence(xr, z)
        loss.backward()
        x64 = x.double().requires_grad_(True)
        loss64 = div.compute_pairwise_divergence(x64, z.double())
        loss64.backward()
        out[f"{name}_x"] = x.numpy()
        out[f"{name}_z"] = z.numpy()
        out[f"{name}_loss"] = loss.detach().numpy()
        out[f"{name}_grad"] = xr.grad.numpy()
        out[f"{name}_loss64"] = loss64.detach().numpy()
        out[f"{name}_grad64"] = x64.grad.numpy()
        if k <= 64:
            out[f"{name}_pairwise"] = div.compute_pairwise(x).numpy()
            out[f"{name}_pairdist"] = div.compute_pair_distance(x).numpy()
        # restatement must reproduce the reference bit for bit (same torch ops)
        l2, g2 = O.ndiv_fwd_bwd(x, z)
        assert torch.equal(l2, loss.detach()) and torch.equal(g2, xr.grad), name
    np.savez_compressed(os.path.join(OUT, "ndiv.npz"), **out)
    print("ndiv.npz", len(out), "arrays")


def golden_gen(RefDecoder):
    out = {}
    for nd in (2, 16):
        torch.manual_seed(0)
        ref = RefDecoder(noise_dim=nd)
        torch.manual_seed(0)
        mine = synthetic.GeneratorMLP(nd)
        for (n1, p1), (n2, p2) in zip(ref.state_dict().items(), mine.state_dict().items()):
            assert n1 == n2 and torch.equal(p1, p2), (n1, n2)
        torch.manual_seed(5)
        zin = torch.randn(33, 256 + nd)
        out[f"nd{nd}_z"] = zin.numpy()
        with torch.no_grad():
            out[f"nd{nd}_a"] = ref.eval()(zin).numpy()
    np.savez_compressed(os.path.join(OUT, "gen.npz"), **out)
    print("gen.npz ok (GeneratorMLP default init == reference Decoder init)")


class FakeDataset(torch.utils.data.Dataset):
    """Duck-typed PushDataset (utils/trajectory_loader.py:38-72) in embedding space."""

    def __init__(self, images, actions):
        self.images, self.actions = images, actions
        self.seq_length = images.shape[1]

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return (self.images[i][:, :, None, None], torch.zeros(self.seq_length, 25),
                self.actions[i], torch.zeros(3))


def golden_mpc(mpc_eval, RefDecoder):
    out = {}
    E, H, nd = 128, 256, 2
    for name, (K, Th, seq, n_traj) in {"small": (64, 3, 4, 2), "cfg1": (1024, 8, 3, 1)}.items():
        gen, fwd = synthetic.make_models(nd=nd, E=E, H=H, seed=0)
        torch.manual_seed(0)
        ref_gen = RefDecoder(noise_dim=nd)
        ref_gen.load_state_dict(gen.state_dict())
        torch.manual_seed(77)
        images = torch.randn(n_traj, seq, E)
        actions = torch.randn(n_traj, seq, 4)
        cfg = DotMap(random_seed=3, gpu_id="cpu", log_port=0,
                     evaluation=dict(num_sample=1, noise_dim=nd, batch_size=1, threshold=0.05),
                     mpc=dict(rollouts=K, time_horizon=Th))
        logging.disable(logging.CRITICAL)
        with contextlib.redirect_stdout(io.StringIO()):
            a_err, i_err = mpc_eval.fetch_push_control_evaluation(
                nn.Identity(), fwd, ref_gen.eval(), FakeDataset(images, actions), cfg)
        logging.disable(logging.NOTSET)
        # restated loop must reproduce the unmodified function's tuple
        a2, i2, best = O.closed_loop_eval(gen, fwd, images, actions, K, Th, nd, seed=3)
        assert abs(a2 - a_err) <= 1e-6 * abs(a_err) and abs(i2 - i_err) <= 1e-6 * abs(i_err), \
            (a_err, a2, i_err, i2)
        print(f"mpc {name}: reference tuple ({a_err:.7g}, {i_err:.7g}) restated ({a2:.7g}, {i2:.7g})")
        out[f"{name}_images"] = images.numpy()
        out[f"{name}_actions"] = actions.numpy()
        out[f"{name}_tuple"] = np.array([a_err, i_err], dtype=np.float64)
        out[f"{name}_best"] = np.array(best, dtype=np.int64)
        out[f"{name}_params"] = np.array([K, Th, seq, n_traj, E, H, nd, 3], dtype=np.int64)
        # one explicit plan with costs (reference modules, restated loop)
        e0, goal, noise = synthetic.make_plan_inputs(1, K, Th, nd, E, seed=21)
        plan = O.mpc_plan(ref_gen, fwd, e0[0], goal[0], noise[:, 0], topk=8)
        out[f"{name}_plan_cost"] = plan["cost"].numpy()
        out[f"{name}_plan_best"] = np.array(plan["best_idx"])
        out[f"{name}_plan_a0"] = plan["a0"].numpy()
        out[f"{name}_plan_topk"] = plan["topk_idx"].numpy()
        g64, f64 = gen.double(), fwd.double()
        plan64 = O.mpc_plan(g64, f64, e0[0].double(), goal[0].double(), noise[:, 0].double(), topk=8)
        out[f"{name}_plan_cost64"] = plan64["cost"].numpy()
    np.savez_compressed(os.path.join(OUT, "mpc.npz"), **out)
    print("mpc.npz", len(out), "arrays")


GAN_STRIDE64 = 4


def flat_params(m):
    return torch.cat([p.detach().reshape(-1) for p in m.parameters()])


def flat_grads(m):
    return torch.cat([p.grad.detach().reshape(-1) for p in m.parameters()])


def golden_gan(div, RefDecoder, RefDiscriminator):
    """train_gan.py:126-203 with the reference's OWN Decoder / Discriminator / diversity modules and
    torch.optim.Adam as constructed at :98-104; the restated step (oracle.gan_train_step) is the loop body."""
    out = {}
    cases = {"ref_shape": (56, 6, 2, 1, 3, 31, 0.1),      # batch 8 x (seq 8 - 1), num_sample 6 (config/default.yaml)
             "s16": (40, 16, 2, 2, 2, 32, 0.1),            # two discriminator steps per generator step
             "nd4": (9, 5, 4, 1, 2, 33, 0.1),              # ragged sizes, wider noise
             "nodiv": (56, 6, 2, 1, 5, 34, 0.0)}           # pairwise_div_factor 0: the smooth part of the recurrence
    for name, (Fb, S, nd, dsteps, nsteps, seed, divf) in cases.items():
        gen_p, dis_p = synthetic.make_gan_models(nd, seed=0)
        results = {}
        for dt in (torch.float32, torch.float64):
            torch.manual_seed(0)
            gen, dis = RefDecoder(noise_dim=nd), RefDiscriminator()
            gen.weight_init(mean=0.0, std=0.02); dis.weight_init(mean=0.0, std=0.02)   # train_gan.py:91-92 (no-ops)
            for (n1, p1), (n2, p2) in zip(gen.state_dict().items(), gen_p.state_dict().items()):
                assert n1 == n2 and torch.equal(p1, p2)
            for (n1, p1), (n2, p2) in zip(dis.state_dict().items(), dis_p.state_dict().items()):
                assert n1 == n2 and torch.equal(p1, p2)
            gen, dis = gen.to(dt).train(), dis.to(dt).train()
            G_opt, D_opt = O.gan_optimizers(gen, dis, lr=2e-4)
            losses = []
            for it in range(nsteps):
                codes, actions, noise = synthetic.make_gan_batch(Fb, S, nd, seed=seed + 100 * it)
                probe = {}
                dl, gl, pl, a_hat = O.gan_train_step(gen, dis, G_opt, D_opt, codes.to(dt), actions.to(dt), noise.to(dt),
                                                     div_factor=divf, discrim_steps_per_gen=dsteps,
                                                     divergence=div.compute_pairwise_divergence, probe=probe)
                losses.append([float(dl), float(gl), float(pl)])
                if it == 0:
                    results[dt, "a_hat0"] = a_hat.numpy()
                    results[dt, "ggrad0"] = flat_grads(gen).numpy()
                    results[dt, "dgrad0"] = probe["d_grad"].numpy()
            results[dt, "losses"] = np.array(losses, dtype=np.float64)
            results[dt, "gparams"] = flat_params(gen).numpy()
            results[dt, "dparams"] = flat_params(dis).numpy()
        # keep the fixture small: parameters every 4th element (GAN_STRIDE64) except the fp32 ones of the reference
        # shape, first-step gradients only for the reference shape
        for (dt, key), v in results.items():
            is64 = dt == torch.float64
            if key in ("ggrad0", "dgrad0") and name != "ref_shape":
                continue
            if key in ("gparams", "dparams") and (is64 or name != "ref_shape"):
                v = v[::GAN_STRIDE64]
            out[f"{name}_{key}{'64' if is64 else ''}"] = v
        out[f"{name}_cfg"] = np.array([Fb, S, nd, dsteps, nsteps, seed], dtype=np.int64)
        out[f"{name}_divf"] = np.array(divf, dtype=np.float64)
        print(f"gan {name}: losses {results[torch.float32, 'losses'][-1]}")
    np.savez_compressed(os.path.join(OUT, "gan.npz"), **out)
    print("gan.npz", len(out), "arrays")


def golden_enc():
    """models/image_autoencoder.py Encoder in eval mode, fp32 and fp64, on seeded frames.  The weights (13 M floats)
    are not committed: synthetic.make_image_encoder(seed) rebuilds them; a checksum guards the RNG stream."""
    from models.image_autoencoder import Encoder as RefEncoder  # type: ignore
    out = {}
    for name, (B, seed) in {"b1": (1, 41), "b2": (2, 42), "b5": (5, 43)}.items():
        enc_p = synthetic.make_image_encoder(seed=7)
        ref = RefEncoder()
        missing = ref.load_state_dict(enc_p.state_dict(), strict=True)          # same names, same shapes
        assert not missing.missing_keys and not missing.unexpected_keys
        assert list(ref.state_dict().keys()) == list(enc_p.state_dict().keys())
        ref.eval()
        frames = synthetic.make_frames(B, seed=seed)
        with torch.no_grad():
            y32 = ref(frames)
            y64 = ref.double()(frames.double())
            mine = O.encoder_forward(enc_p.state_dict(), frames.double())
        assert y32.shape == (B, 128, 1, 1)
        err = float((mine - y64).abs().max() / y64.abs().max())
        assert err < 1e-12, err
        out[f"{name}_cfg"] = np.array([B, seed, 7], dtype=np.int64)
        out[f"{name}_codes"] = y32.view(B, 128).numpy()
        out[f"{name}_codes64"] = y64.view(B, 128).numpy()
        out[f"{name}_frame_sum"] = np.array(frames.double().sum().item())
        print(f"enc {name}: |codes| max {float(y64.abs().max()):.4f}, fp32-vs-fp64 {float((y32.double() - y64).abs().max()):.3e}, "
              f"oracle-vs-ref(fp64) {err:.2e}")
    # default-initialised container draws the same parameters as the reference class under the same seed
    from ndivplanning_b200.models import ImageEncoder
    torch.manual_seed(3); a = RefEncoder()
    torch.manual_seed(3); b = ImageEncoder()
    for (n1, p1), (n2, p2) in zip(a.state_dict().items(), b.state_dict().items()):
        assert n1 == n2 and torch.equal(p1, p2), n1
    out["param_sum"] = np.array(sum(float(v.double().sum()) for v in synthetic.make_image_encoder(seed=7).state_dict().values()))
    np.savez_compressed(os.path.join(OUT, "enc.npz"), **out)
    print("enc.npz", len(out), "arrays")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    div, mpc_eval, RefDecoder = import_reference()
    golden_ndiv(div)
    golden_gen(RefDecoder)
    golden_mpc(mpc_eval, RefDecoder)
    from models.gan import Discriminator as RefDiscriminator  # type: ignore  # noqa: E402
    golden_gan(div, RefDecoder, RefDiscriminator)
    golden_enc()
