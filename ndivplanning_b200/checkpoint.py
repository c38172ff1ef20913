"""Checkpoint ingestion (the part that needs no data files): the reference stores WHOLE pickled
modules with torch.save(model, path) and reloads them with torch.load(path, map_location=gpu_id) (mpc_eval.py:214-219,
train_gan.py:82-92).  These helpers accept such files — or plain state_dict files, or live modules — and hand the
weights to the planner / trainer; only parameter names and shapes matter (fc1..fc5 of models.gan.Decoder,
fc1..fc4 of models.gan.Discriminator, fc1/fc2 of the latent forward model), so a pickle of the reference's own
classes works wherever its defining module is importable.  No compute happens here.
"""
from typing import Union

import torch
from torch import nn

from . import _lib
from .models import ActionDiscriminator, GeneratorMLP, LatentForward

Source = Union[str, nn.Module, dict]


def _state_dict(src: Source, trust_pickle: bool = False) -> dict:
    if isinstance(src, (str, bytes)) or hasattr(src, "__fspath__"):
        try:
            src = torch.load(src, map_location="cpu", weights_only=True)  # state_dict files: no code execution
        except Exception as e:
            # the reference saves WHOLE modules (torch.save(model, path)): unpickling those executes code from the
            # file, so it only happens on explicit request
            if not trust_pickle:
                raise _lib.NdpError(f"{src}: not a plain state_dict file ({type(e).__name__}); whole pickled modules "
                                    f"(the reference's torch.save(model)) need trust_pickle=True") from e
            src = torch.load(src, map_location="cpu", weights_only=False)
    if isinstance(src, nn.Module):
        return src.state_dict()
    if isinstance(src, dict):
        return src["state_dict"] if "state_dict" in src and isinstance(src["state_dict"], dict) else src
    raise _lib.NdpError(f"cannot read weights from {type(src).__name__}")


def load_generator(src: Source, trust_pickle: bool = False) -> GeneratorMLP:
    sd = _state_dict(src, trust_pickle)
    if "fc1.weight" not in sd or "fc5.weight" not in sd:
        raise _lib.NdpError("generator checkpoint does not have the models.gan.Decoder layout (fc1..fc5)")
    cond = 256
    m = GeneratorMLP(sd["fc1.weight"].shape[1] - cond, cond_dim=cond)
    m.load_state_dict({k: v for k, v in sd.items() if k.split(".")[0] in ("fc1", "fc2", "fc3", "fc4", "fc5")})
    return m.eval()


def load_discriminator(src: Source, trust_pickle: bool = False) -> ActionDiscriminator:
    sd = _state_dict(src, trust_pickle)
    if "fc4.weight" not in sd or sd["fc4.weight"].shape[0] != 1:
        raise _lib.NdpError("discriminator checkpoint does not have the models.gan.Discriminator layout (fc1..fc4)")
    m = ActionDiscriminator(cond_dim=sd["fc1.weight"].shape[1] - 4)
    m.load_state_dict({k: v for k, v in sd.items() if k.split(".")[0] in ("fc1", "fc2", "fc3", "fc4")})
    return m


def load_forward_model(src: Source, trust_pickle: bool = False) -> LatentForward:
    sd = _state_dict(src, trust_pickle)
    if "fc1.weight" not in sd or "fc2.weight" not in sd:
        raise _lib.NdpError("forward-model checkpoint does not have the LatentForward layout (fc1, fc2)")
    H, EA = sd["fc1.weight"].shape
    m = LatentForward(E=sd["fc2.weight"].shape[0], H=H, A=EA - sd["fc2.weight"].shape[0])
    m.load_state_dict({k: v for k, v in sd.items() if k.split(".")[0] in ("fc1", "fc2")})
    return m.eval()


def load_planner(generator: Source, fwd_model: Source, *, horizon: int, precision: str = "f16+refine", device=None,
                 trust_pickle: bool = False):
    """mpc_eval.py:214-219 -> a FusedPlanner (weights packed once on the device)"""
    from .planner import FusedPlanner
    return FusedPlanner(load_generator(generator, trust_pickle), load_forward_model(fwd_model, trust_pickle),
                        horizon=horizon, precision=precision, device=device)


def load_trainer(decoder: Source, discriminator: Source, trust_pickle: bool = False, **kw):
    """a GanTrainer resumed from decoder / discriminator checkpoints (optimiser state starts fresh, as in the
    reference, which never saves it)"""
    from .gan_train import GanTrainer
    return GanTrainer(load_generator(decoder, trust_pickle).train(), load_discriminator(discriminator, trust_pickle).train(), **kw)
