"""GPU parity of the device-resident image encoder (first slice; csrc/k5_encoder.cu through
ndp_encoder_forward) against tests/golden/enc.npz — the reference's models/image_autoencoder.py Encoder in eval mode
run by oracle/make_golden.py — and against the oracle restatement.  Floating point: the kernel accumulates each
output in fp32 over up to 16384 terms in a different order than the reference's convolution; the bound is a multiple
of the reference's own fp32-vs-fp64 error (the fp64 run arbitrates), plus an absolute floor written below."""
import numpy as np
import pytest
import torch

from ndivplanning_b200 import _lib, synthetic
from ndivplanning_b200.encoder import FusedEncoder
from oracle import ndiv_oracle as O

pytestmark = pytest.mark.gpu
FLOOR = 2e-6        # absolute, on codes of magnitude ~0.5


@pytest.fixture(scope="module")
def fused(cuda):
    enc = synthetic.make_image_encoder(seed=7)
    return enc, FusedEncoder(enc, max_images=8, device=cuda)


@pytest.mark.parametrize("name", ["b1", "b2", "b5"])
def test_encoder_matches_reference_run(cuda, golden, fused, name):
    g = golden["enc"]
    B, seed, _ = (int(v) for v in g[f"{name}_cfg"])
    frames = synthetic.make_frames(B, seed=seed)
    y = fused[1](frames.to(cuda))
    assert y.shape == (B, 128, 1, 1)
    mine = y.view(B, 128).cpu().numpy().astype(np.float64)
    ref64, ref32 = g[f"{name}_codes64"], g[f"{name}_codes"].astype(np.float64)
    e_m, e_r = np.abs(mine - ref64), np.abs(ref32 - ref64)
    assert e_m.max() <= 4 * e_r.max() + FLOOR, (e_m.max(), e_r.max())
    assert np.sqrt((e_m ** 2).mean()) <= 4 * np.sqrt((e_r ** 2).mean()) + FLOOR / 4


def test_encoder_chunks_and_determinism(cuda, fused):
    """B > max_images goes through in chunks; every image's code is independent of its batch (split-K plan changes
    with B, so this is a tolerance check) and a repeat of the same call is bit-identical."""
    enc, fe = fused
    frames = synthetic.make_frames(11, seed=9).to(cuda)
    a = fe(frames)
    b = fe(frames)
    assert torch.equal(a, b)
    one = torch.cat([fe(frames[i:i + 1]) for i in range(11)])
    assert float((a - one).abs().max()) < FLOOR
    with torch.no_grad():
        want = O.encoder_forward(enc.state_dict(), frames.cpu().double()).view(11, 128)
    assert float((a.view(11, 128).cpu().double() - want).abs().max()) < 4 * FLOOR
    assert fe(frames[:0]).shape == (0, 128, 1, 1)


def test_encoder_untrained_default_init(cuda):
    """Default-initialised parameters (BatchNorm at identity statistics) and black / white frames."""
    from ndivplanning_b200.models import ImageEncoder
    torch.manual_seed(5)
    enc = ImageEncoder().eval()
    fe = FusedEncoder(enc, max_images=4, device=cuda)
    frames = torch.stack([torch.zeros(3, 128, 128), torch.ones(3, 128, 128), synthetic.make_frames(1, seed=2)[0]])
    with torch.no_grad():
        want = O.encoder_forward(enc.state_dict(), frames.double()).view(3, 128)
    got = fe(frames.to(cuda)).view(3, 128).cpu().double()
    assert float((got - want).abs().max()) < FLOOR


def test_encoder_refuses(cuda, fused):
    enc, fe = fused
    with pytest.raises(_lib.NdpError):
        fe(torch.zeros(1, 3, 128, 128))                    # host tensor: no CPU path
    with pytest.raises(ValueError):
        fe(torch.zeros(1, 3, 64, 64, device=cuda))
    with pytest.raises(ValueError):
        FusedEncoder(synthetic.make_image_encoder(seed=1).train(), device=cuda)
    with pytest.raises(_lib.NdpError):
        h = __import__("ctypes").c_void_p()
        _lib.check(_lib.lib().ndp_encoder_create(__import__("ctypes").byref(h), 4, 96))


def test_encoder_feeds_planner_codes(cuda, fused):
    """mpc_eval.py:135-136: the planner's e0 / goal are encoder outputs flattened to [G,128]."""
    enc, fe = fused
    frames = synthetic.make_frames(2, seed=21).to(cuda)
    codes = fe(frames).flatten(1)
    assert codes.shape == (2, 128) and codes.is_contiguous() and torch.isfinite(codes).all()


def test_closed_loop_from_frames_matches_oracle(cuda):
    """fetch_push_control_evaluation in frame mode: dataset images are camera frames, encoded by the device encoder;
    the oracle runs the same receding-horizon loop (mpc_eval.py:112-176) on the oracle's codes."""
    from ndivplanning_b200 import fetch_push_control_evaluation
    from oracle.ref_stubs import DotMap
    K, Th, seq, n_traj, nd, seed = 64, 3, 4, 2, 2, 13
    enc = synthetic.make_image_encoder(seed=7)
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=128, H=256, seed=0)
    frames = synthetic.make_frames(n_traj * seq, seed=17).view(n_traj, seq, 3, 128, 128)
    actions = torch.rand(n_traj, seq, 4, generator=torch.Generator().manual_seed(3))
    with torch.no_grad():
        codes = O.encoder_forward(enc.state_dict(), frames.flatten(0, 1)).view(n_traj, seq, 128)
        want_a, want_i, want_best = O.closed_loop_eval(gen_p, fwd_p, codes, actions, K, Th, nd, seed)

    class DS(torch.utils.data.Dataset):
        seq_length = seq
        def __len__(self): return n_traj
        def __getitem__(self, i):
            return frames[i], torch.zeros(seq, 25), actions[i], torch.zeros(3)

    cfg = DotMap(random_seed=seed, gpu_id=0, log_port=0,
                 evaluation=dict(num_sample=1, noise_dim=nd, batch_size=1, threshold=0.05),
                 mpc=dict(rollouts=K, time_horizon=Th))
    a, i, best = fetch_push_control_evaluation(enc, fwd_p, gen_p, DS(), cfg, return_indices=True)
    assert best.numpy().tolist() == [list(r) for r in want_best]
    np.testing.assert_allclose([a, i], [want_a, want_i], rtol=1e-4)


def test_closed_loop_from_push_dataset_bundles(cuda, tmp_path):
    """the same loop fed by the PushDataset drop-in reading trajectory bundles from disk (decoded uint8 frames ->
    the reference's [-1,1] normalisation -> device encoder): equals the oracle loop on the oracle's codes of exactly
    those normalised frames"""
    from ndivplanning_b200 import PushDataset, fetch_push_control_evaluation, trajectory_loader
    from oracle.ref_stubs import DotMap
    K, Th, seq, n_traj, nd, seed = 64, 3, 4, 3, 2, 29
    enc = synthetic.make_image_encoder(seed=7)
    gen_p, fwd_p = synthetic.make_models(nd=nd, E=128, H=256, seed=0)
    rng = np.random.default_rng(5)
    base = ((synthetic.make_frames(n_traj * seq, seed=23) * 0.5 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)
    u8 = base.view(n_traj, seq, 3, 128, 128).permute(0, 1, 3, 4, 2).contiguous().numpy()
    acts = rng.random((n_traj, seq, 4)).astype(np.float32)
    trajs = [(u8[i], np.zeros((seq, 25), np.float32), acts[i], np.zeros(3, np.float32)) for i in range(n_traj)]
    trajectory_loader.write_npz_bundle(str(tmp_path / "trajectory_bundle_00000.npz"), trajs[:2])
    trajectory_loader.write_npz_bundle(str(tmp_path / "trajectory_bundle_00001.npz"), trajs[2:])
    ds = PushDataset(str(tmp_path), seq_length=seq)
    frames = torch.stack([ds[i][0] for i in range(n_traj)])
    with torch.no_grad():
        codes = O.encoder_forward(enc.state_dict(), frames.flatten(0, 1)).view(n_traj, seq, 128)
        want_a, want_i, want_best = O.closed_loop_eval(gen_p, fwd_p, codes, torch.from_numpy(acts), K, Th, nd, seed)
    cfg = DotMap(random_seed=seed, gpu_id=0, log_port=0,
                 evaluation=dict(num_sample=1, noise_dim=nd, batch_size=1, threshold=0.05),
                 mpc=dict(rollouts=K, time_horizon=Th))
    a, i, best = fetch_push_control_evaluation(enc, fwd_p, gen_p, ds, cfg, return_indices=True)
    assert best.numpy().tolist() == [list(r) for r in want_best]
    np.testing.assert_allclose([a, i], [want_a, want_i], rtol=1e-4)


def test_gan_epoch_from_frames_matches_oracle(cuda, fused):
    """train_gan.py:114-209 with a real image encoder in front of the step (:152-155): FusedEncoder feeds
    GanTrainer on the device; the oracle epoch encodes with oracle.encoder_forward on the CPU."""
    from ndivplanning_b200 import GanTrainer, train_gan_epoch
    enc, fe = fused
    seq, bs, n, S, nd = 4, 2, 4, 6, 2
    frames = synthetic.make_frames(n * seq, seed=23).view(n, seq, 3, 128, 128)
    actions = torch.randn(n, seq, 4, generator=torch.Generator().manual_seed(1))

    class DS(torch.utils.data.Dataset):
        seq_length = seq
        def __len__(self): return n
        def __getitem__(self, i):
            return frames[i], torch.zeros(seq, 25), actions[i], torch.zeros(3)

    loader = torch.utils.data.DataLoader(DS(), batch_size=bs, shuffle=False)
    F = bs * (seq - 1)
    torch.manual_seed(3)
    noises = [torch.rand(F, S, nd) for _ in range(len(loader))]
    gen, dis = synthetic.make_gan_models(nd, seed=0)
    tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=F, pairwise_div_factor=0.0, device=cuda)
    it = iter(noises)
    got = train_gan_epoch(tr, loader, fe, seq, noise_fn=lambda F_: next(it))
    og, od = O.GeneratorMLP(nd), O.ActionDiscriminator()
    og.load_state_dict(gen.state_dict()); od.load_state_dict(dis.state_dict())
    G_opt, D_opt = O.gan_optimizers(og, od)
    sd = enc.state_dict()
    want = O.gan_train_epoch(og, od, G_opt, D_opt, loader, lambda x: O.encoder_forward(sd, x), seq, S, nd,
                             div_factor=0.0, noises=noises)
    np.testing.assert_allclose(got[:2], want[:2], rtol=2e-5)
    np.testing.assert_allclose(got[2], want[2], rtol=5e-3)


def test_encoder_large_batch_kernels(cuda):
    """B = 32 and a ragged B = 27 take the 128x128-tile kernel (fused bias + ReLU epilogue, split-K on the deep
    layers); same bound against the fp64 oracle."""
    enc = synthetic.make_image_encoder(seed=7)
    fe = FusedEncoder(enc, max_images=32, device=cuda)
    frames = synthetic.make_frames(32, seed=31)
    with torch.no_grad():
        want = O.encoder_forward(enc.state_dict(), frames.double()).view(32, 128)
    got = fe(frames.to(cuda)).view(32, 128).cpu().double()
    assert float((got - want).abs().max()) < 4 * FLOOR
    got27 = fe(frames[:27].to(cuda)).view(27, 128).cpu().double()
    assert float((got27 - want[:27]).abs().max()) < 4 * FLOOR


def test_encoder_cuda_graph_replay(cuda, fused):
    """the forward is capturable: replays over refilled static frames are bit-identical to eager calls"""
    enc, fe = fused
    frames = synthetic.make_frames(8, seed=51).to(cuda)
    static = frames[:2].clone()
    g, codes = fe.capture(static)
    for lo in (0, 2, 4):
        static.copy_(frames[lo:lo + 2])
        g.replay()
        assert torch.equal(codes, fe(frames[lo:lo + 2]))
