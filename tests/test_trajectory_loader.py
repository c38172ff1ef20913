"""Host-side parity of the PushDataset drop-in (reference utils/trajectory_loader.py:17-72,
utils/hdf5_load.py:9-11): index -> (file, trajectory) mapping over several bundles, seq_start / seq_length slicing,
the reference's frame normalisation, the JPEG path, the h5py-module hook, and the loud failure without h5py."""
import io
import os
import types

import numpy as np
import pytest
import torch

from ndivplanning_b200 import trajectory_loader as TL


def _traj(rng, S=6, H=16, W=16):
    return (rng.integers(0, 256, (S, H, W, 3), dtype=np.uint8), rng.standard_normal((S, 25)).astype(np.float32),
            rng.standard_normal((S, 4)).astype(np.float32), rng.standard_normal(3).astype(np.float32))


def test_npz_bundles_index_mapping_and_normalisation(tmp_path):
    rng = np.random.default_rng(0)
    a, b = [_traj(rng) for _ in range(3)], [_traj(rng) for _ in range(2)]
    TL.write_npz_bundle(str(tmp_path / "trajectory_bundle_00000.npz"), a)
    TL.write_npz_bundle(str(tmp_path / "trajectory_bundle_00001.npz"), b)
    ds = TL.PushDataset(str(tmp_path), seq_start=1, seq_length=4)
    assert len(ds) == 5 and ds.seq_length == 4
    for i, (fr, st, ac, goal) in enumerate(a + b):
        images, states, actions, g = ds[i]
        assert images.shape == (4, 3, 16, 16) and images.dtype == torch.float32
        want = (torch.from_numpy(fr[1:5]).permute(0, 3, 1, 2).float() / 255.0 - 0.5) * 2.0   # hdf5_load.py:11
        assert torch.equal(images, want)
        assert torch.equal(states, torch.from_numpy(st[1:5])) and torch.equal(actions, torch.from_numpy(ac[1:5]))
        assert torch.equal(g, torch.from_numpy(goal))
    assert torch.equal(ds[-1][3], torch.from_numpy(b[-1][3]))
    with pytest.raises(IndexError):
        ds[5]
    tf = TL.PushDataset(str(tmp_path), seq_length=2, transform=lambda x: x * 0)
    assert float(tf[0][0].abs().max()) == 0.0


def test_jpeg_frames_decode_like_the_reference(tmp_path):
    from PIL import Image
    rng = np.random.default_rng(1)
    t = [_traj(rng, S=3, H=32, W=32)]
    TL.write_npz_bundle(str(tmp_path / "b.npz"), t, jpeg_quality=90)
    ds = TL.PushDataset(str(tmp_path), seq_length=3)
    images = ds[0][0]
    assert images.shape == (3, 3, 32, 32) and float(images.min()) >= -1.0 and float(images.max()) <= 1.0
    raw = np.load(str(tmp_path / "b.npz"), allow_pickle=True)["trajectory_00000/images"]
    for s in range(3):                                       # the reference's own decode chain, spelled out
        arr = np.asarray(Image.open(io.BytesIO(raw[s])), dtype=np.uint8)
        want = (torch.from_numpy(arr.copy()).permute(2, 0, 1).float() / 255.0 - 0.5) * 2.0
        assert torch.equal(images[s], want)
        assert torch.equal(TL.image_bytes_to_norm_tensor(raw[s]), want)


def test_h5_module_hook_and_loud_failure(tmp_path):
    rng = np.random.default_rng(2)
    t = [_traj(rng) for _ in range(2)]
    TL.write_npz_bundle(str(tmp_path / "store.npz"), t)
    store = str(tmp_path / "store.npz")
    h5dir = tmp_path / "h5"
    h5dir.mkdir()
    (h5dir / "trajectory_bundle_00000.h5").write_bytes(b"\x89HDF\r\n\x1a\n")
    calls = []
    fake = types.SimpleNamespace(File=lambda path, mode: (calls.append((os.path.basename(path), mode)), TL._NpzBundle(store))[1])
    ds = TL.PushDataset(str(h5dir), seq_length=5, h5_module=fake)
    assert len(ds) == 2 and torch.equal(ds[1][2], torch.from_numpy(t[1][2][:5]))
    assert calls and all(c == ("trajectory_bundle_00000.h5", "r") for c in calls)
    try:
        import h5py  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError, match="h5py"):
            TL.PushDataset(str(h5dir))
    with pytest.raises(FileNotFoundError):
        TL.PushDataset(str(tmp_path / "h5" / "nothing_here"))
