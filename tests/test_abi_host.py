"""CPU: the C-ABI library loads, exports every symbol include/ndivplan.h declares, refuses to
compute without a GPU; the product never touches oracle/."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "ndivplan.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ndp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from ndivplanning_b200 import _lib
    lib = _lib.lib()
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in ndivplan.h but not exported"
    assert set(_lib.EXPORTED_SYMBOLS) == set(declared)
    assert lib.ndp_abi_version() == 1


def test_no_cpu_fallback():
    from ndivplanning_b200 import NdpError, diversity, select_plan
    x, z = torch.randn(2, 4, 3), torch.rand(2, 4, 2)
    with pytest.raises(NdpError):
        diversity.compute_pairwise_divergence(x, z)
    with pytest.raises(NdpError):
        select_plan(torch.rand(1, 8))
    if not torch.cuda.is_available():
        from ndivplanning_b200 import _lib
        assert _lib.lib().ndp_device_check() != 0
        # argument validation happens before any device work
        n = ctypes.c_size_t()
        assert _lib.lib().ndp_select_workspace_bytes(1, 10, 500, ctypes.byref(n)) == -1
        assert b"topk" in _lib.lib().ndp_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "ndivplanning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "/root/reference" not in txt, f


def test_synthetic_spec_matches_oracle_modules():
    from ndivplanning_b200 import synthetic
    from oracle import ndiv_oracle as O
    gen, fwd = synthetic.make_models(nd=2, E=128, H=256)
    torch.manual_seed(0)
    og, of = O.GeneratorMLP(2), O.LatentForward(128, 256)
    with torch.no_grad():
        og.fc1.weight[:, 256:] *= 30
        of.fc1.weight[:, 128:] *= 30
    for a, b in zip(list(gen.parameters()) + list(fwd.parameters()), list(og.parameters()) + list(of.parameters())):
        assert torch.equal(a, b)
    e0, goal, noise = synthetic.make_plan_inputs(2, 16, 3)
    assert noise.shape == (3, 2, 16, 2) and float(noise.min()) >= 0 and float(noise.max()) < 1


def test_checkpoint_ingestion_roundtrip(tmp_path):
    """whole pickled modules (how the reference saves: torch.save(model, path)), state_dict files and wrapped dicts
    all load into the weight containers the planner / trainer read (CPU only: no compute)"""
    import pytest
    import torch
    from ndivplanning_b200 import checkpoint, synthetic
    gen, fwd = synthetic.make_models(nd=3, E=128, H=256)
    _, dis = synthetic.make_gan_models(3, seed=1)
    pg, pf, pd = tmp_path / "gen.pt", tmp_path / "fwd.pt", tmp_path / "dis.pt"
    torch.save(gen, pg)                                   # whole module
    torch.save(fwd.state_dict(), pf)                      # plain state_dict
    torch.save({"state_dict": dis.state_dict(), "epoch": 7}, pd)
    with pytest.raises(Exception):
        checkpoint.load_generator(str(pg))               # whole pickled module: only on explicit trust
    g2 = checkpoint.load_generator(str(pg), trust_pickle=True)
    f2, d2 = checkpoint.load_forward_model(pf), checkpoint.load_discriminator(pd)
    for a, b in ((gen, g2), (fwd, f2), (dis, d2)):
        for (k1, v1), (k2, v2) in zip(a.state_dict().items(), b.state_dict().items()):
            assert k1 == k2 and torch.equal(v1, v2)
    assert g2.fc1.weight.shape[1] == 256 + 3 and f2.fc1.weight.shape == (256, 132)
    with pytest.raises(Exception):
        checkpoint.load_discriminator(str(pg), trust_pickle=True)      # a generator is not a discriminator
