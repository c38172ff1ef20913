"""Drop-in for the reference's trajectory dataset (reference utils/trajectory_loader.py:17-72,
utils/hdf5_load.py:9-11): ``PushDataset(datadir, seq_start=0, seq_length=15, transform=None)`` with
``dataset[i] -> (images [S,3,H,W] in [-1,1], states [S,25], actions [S,4], goal [3])`` — what
``fetch_push_control_evaluation`` / ``train_gan_epoch`` iterate over.

On-disk format (generate_trajectories.py:259-324): one HDF5 file per bundle, one group ``trajectory_%05d`` per
trajectory holding ``images`` (one JPEG byte string per frame), ``states``, ``actions``, ``goal``.  Reading HDF5 needs
h5py, which is imported lazily: where it is absent (this image) the HDF5 backend fails loudly.  The same class also
reads ``.npz`` bundles with the same keys (``trajectory_00000/images`` ... as object / uint8 arrays of JPEG bytes, or
already decoded ``uint8 [S,H,W,3]`` frames), which is what the tests and ``write_npz_bundle`` use.  Host-side IO only —
nothing here touches the GPU; frames are decoded with PIL exactly as the reference does
(``(ToTensor()(Image.open(bytes)) - 0.5) * 2``: uint8 HWC -> float CHW / 255, then [-1, 1])."""
import glob
import io
import os

import numpy as np
import torch


def image_bytes_to_norm_tensor(raw_bytes) -> torch.Tensor:
    """utils/hdf5_load.py:9-11 without torchvision: JPEG bytes -> float32 [3,H,W] in [-1,1]"""
    from PIL import Image
    img = Image.open(io.BytesIO(bytes(raw_bytes)))
    arr = np.asarray(img.convert("RGB") if img.mode != "RGB" else img, dtype=np.uint8)
    t = torch.from_numpy(arr.copy()).permute(2, 0, 1).to(torch.float32) / 255.0       # ToTensor()
    return (t - 0.5) * 2.0


def _frames_to_norm_tensor(frames) -> torch.Tensor:
    """decoded uint8 frames [S,H,W,3] -> float32 [S,3,H,W] in [-1,1] (same arithmetic as the JPEG path)"""
    t = torch.from_numpy(np.ascontiguousarray(frames)).permute(0, 3, 1, 2).to(torch.float32) / 255.0
    return (t - 0.5) * 2.0


class _NpzBundle:
    """read-only view of an .npz bundle with h5py-like access: len(), bundle["trajectory_00003"]["images"]"""

    def __init__(self, path):
        self._z = np.load(path, allow_pickle=True)
        self._groups = sorted({k.split("/", 1)[0] for k in self._z.files})

    def __len__(self):
        return len(self._groups)

    def __getitem__(self, group):
        if group not in self._groups:
            raise KeyError(group)
        z = self._z
        return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(group + "/")}

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self._z.close()


def _open_bundle(path, h5_module=None):
    if path.endswith(".npz"):
        return _NpzBundle(path)
    if h5_module is None:
        try:
            import h5py as h5_module                                            # noqa: N813
        except ImportError as e:
            raise ImportError(f"{path}: reading HDF5 trajectory bundles needs h5py, which is not installed; convert "
                              f"the bundles to .npz (trajectory_loader.write_npz_bundle) or install h5py") from e
    return h5_module.File(path, "r")


class PushDataset(torch.utils.data.Dataset):
    """utils/trajectory_loader.py:17-72.  `h5_module`: an h5py-compatible module (tests inject a stand-in)."""

    def __init__(self, datadir, seq_start=0, seq_length=15, transform=None, h5_module=None):
        self.datadir = datadir
        # the reference lists with glob (arbitrary order, trajectory_loader.py:20 / argparse_util.py:22-23); sorted
        # here so that an index names the same trajectory on every machine
        self.files = sorted(glob.glob(os.path.join(datadir, "*")))
        if not self.files:
            raise FileNotFoundError(f"PushDataset: no trajectory bundles under {datadir!r}")
        self.transform = transform
        self.seq_start = seq_start
        self.seq_length = seq_length
        self._h5 = h5_module
        counts = []
        for file in self.files:
            with _open_bundle(file, h5_module) as f:
                counts.append(len(f))
        self.total_seq_ct = int(sum(counts))
        self.file_seq_cts = np.cumsum(counts)

    def __len__(self):
        return self.total_seq_ct

    def __getitem__(self, index):
        if index < 0:
            index += self.total_seq_ct
        if not 0 <= index < self.total_seq_ct:
            raise IndexError(index)
        file_index = int(np.argmax(self.file_seq_cts > index))                 # trajectory_loader.py:40-45
        seq_index = index if file_index == 0 else index - int(self.file_seq_cts[file_index - 1])
        lo, hi = self.seq_start, self.seq_start + self.seq_length
        with _open_bundle(self.files[file_index], self._h5) as f:
            seq = f["trajectory_{:05d}".format(seq_index)]
            raw = seq["images"][lo:hi]
            if isinstance(raw, np.ndarray) and raw.dtype == np.uint8 and raw.ndim == 4:
                images = _frames_to_norm_tensor(raw)
            else:
                images = torch.stack([image_bytes_to_norm_tensor(b) for b in raw])
            states = torch.from_numpy(np.asarray(seq["states"][lo:hi]))
            actions = torch.from_numpy(np.asarray(seq["actions"][lo:hi]))
            goal = torch.from_numpy(np.array(seq["goal"]))
            if self.transform:
                images = self.transform(images)
        return images, states, actions, goal


def write_npz_bundle(path, trajectories, jpeg_quality=None):
    """trajectories: iterable of (frames uint8 [S,H,W,3], states [S,25], actions [S,4], goal [3]).  With
    `jpeg_quality` the frames are stored as JPEG byte strings like the reference's HDF5 bundles, else decoded."""
    out = {}
    for i, (frames, states, actions, goal) in enumerate(trajectories):
        g = "trajectory_{:05d}".format(i)
        frames = np.asarray(frames, dtype=np.uint8)
        if jpeg_quality is not None:
            from PIL import Image
            enc = np.empty(len(frames), dtype=object)
            for s, fr in enumerate(frames):
                buf = io.BytesIO()
                Image.fromarray(fr).save(buf, format="JPEG", quality=int(jpeg_quality))
                enc[s] = buf.getvalue()
            out[g + "/images"] = enc
        else:
            out[g + "/images"] = frames
        out[g + "/states"] = np.asarray(states)
        out[g + "/actions"] = np.asarray(actions)
        out[g + "/goal"] = np.asarray(goal)
    np.savez(path, **out)
