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
