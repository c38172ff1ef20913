"""Device-resident GAN training step around the ndiv loss — SURVEY.md 8f row N2; the loop body of the
reference's train_gan.py:159-203 on encoded inputs:

    codes   = cat([encoder(state_cur), encoder(state_target)], 1)      # [F,256], detached (:152-156)
    a_hat   = decoder(cat([codes repeated S times, noise], -1))        # :159-165
    D_loss  = BCE(D(actions, codes), 1) + BCE(D(a_hat, codes), 0); D_optimizer.step()      # :172-184
    G_loss  = BCE(D(a_hat, codes), 1)                                  # :187-190
    div     = diversity.compute_pairwise_divergence(a_hat.view(F,S,-1), noise)             # :193-196
    (G_loss + pairwise_div_factor * div).backward(); G_optimizer.step()                    # :198-203

``GanTrainer`` keeps the reference's objects at the boundary — a ``models.gan.Decoder`` / ``Discriminator`` pair
(or the product's ``GeneratorMLP`` / ``ActionDiscriminator``), the hyper-parameters of ``config.training.gan`` —
and runs every iteration through ``ndp_gan_train_step`` (csrc/k4_gan_train.cu): parameters, Adam state and all
activations stay on the GPU, nothing synchronises with the host, and a step can be replayed as a CUDA graph.
No compute happens in Python and there is no CPU path.
"""
import ctypes as C
from typing import Optional

import torch

from . import _lib
from .encoder import FusedEncoder

G_LAYERS = ("fc1", "fc2", "fc3", "fc4", "fc5")     # models/gan.py:67-71
D_LAYERS = ("fc1", "fc2", "fc3", "fc4")            # models/gan.py:93-96


class GanTrainer:
    def __init__(self, decoder, discriminator, *, num_sample: int, noise_dim: int, max_codes: int,
                 learning_rate: float = 2e-4, betas=(0.5, 0.999), eps: float = 1e-8,
                 pairwise_div_factor: float = 0.1, discrim_steps_per_gen: int = 1, margin: float = 0.8,
                 device=None):
        if not torch.cuda.is_available():
            raise _lib.NdpError("GanTrainer needs an sm_100a GPU; there is no CPU fallback")
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        if self.device.type != "cuda":
            raise _lib.NdpError("GanTrainer needs an sm_100a GPU; there is no CPU fallback")
        self.decoder, self.discriminator = decoder, discriminator
        gsd, dsd = decoder.state_dict(), discriminator.state_dict()
        self.C = gsd["fc1.weight"].shape[1] - noise_dim
        self.A = gsd["fc5.weight"].shape[0]
        if dsd["fc1.weight"].shape[1] != self.C + self.A:
            raise ValueError("discriminator fc1 does not take cat([action, state_code])")
        self.S, self.nd, self.max_F = int(num_sample), int(noise_dim), int(max_codes)
        self.lr = float(learning_rate)
        desc = _lib.GanDesc(self.C, self.nd, self.A, self.S, self.max_F, int(discrim_steps_per_gen), float(margin),
                            float(pairwise_div_factor), self.lr, float(betas[0]), float(betas[1]), float(eps))
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_create(C.byref(self._h), C.byref(desc)))
        self.losses = torch.zeros(3, dtype=torch.float32, device=self.device)
        self._graph = None
        self.load(gsd, dsd)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().ndp_gan_destroy(h)
            except Exception:       # interpreter shutdown
                pass
            self._h = None

    # ---- parameters ----------------------------------------------------------------------------------------
    def _stage(self, sd, layers):
        W = [_lib.f32c(sd[f"{n}.weight"]).to(self.device) for n in layers]
        b = [_lib.f32c(sd[f"{n}.bias"]).to(self.device) for n in layers]
        return W, b

    @staticmethod
    def _ptrs(ts):
        return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])

    def load(self, gen_state_dict, dis_state_dict) -> None:
        """copies the parameters into the handle and resets both optimisers"""
        gW, gb = self._stage(gen_state_dict, G_LAYERS)
        dW, db = self._stage(dis_state_dict, D_LAYERS)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_load(self._h, self._ptrs(gW), self._ptrs(gb), self._ptrs(dW), self._ptrs(db),
                                               _lib.stream_ptr(self.device)))
            torch.cuda.current_stream(self.device).synchronize()      # staged copies may now be freed

    def state_dicts(self):
        """(generator, discriminator) parameters as fresh state dicts (device tensors)"""
        gW, gb = self._stage(self.decoder.state_dict(), G_LAYERS)
        dW, db = self._stage(self.discriminator.state_dict(), D_LAYERS)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_export(self._h, self._ptrs(gW), self._ptrs(gb), self._ptrs(dW), self._ptrs(db),
                                                 _lib.stream_ptr(self.device)))
        g = {f"{n}.{k}": t for n, w, b in zip(G_LAYERS, gW, gb) for k, t in (("weight", w), ("bias", b))}
        d = {f"{n}.{k}": t for n, w, b in zip(D_LAYERS, dW, db) for k, t in (("weight", w), ("bias", b))}
        return g, d

    def sync_modules(self) -> None:
        """writes the trained parameters back into the decoder / discriminator modules"""
        g, d = self.state_dicts()
        self.decoder.load_state_dict({k: v.to(self.decoder.fc1.weight.device) for k, v in g.items()})
        self.discriminator.load_state_dict({k: v.to(self.discriminator.fc1.weight.device) for k, v in d.items()})

    def set_lr(self, lr: float) -> None:
        """StepLR of train_gan.py:105 lives in the caller; the value is read by the next optimiser step"""
        self.lr = float(lr)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_set_lr(self._h, self.lr, _lib.stream_ptr(self.device)))

    # ---- the step --------------------------------------------------------------------------------------------
    def step(self, codes: torch.Tensor, actions: torch.Tensor, noise: torch.Tensor, comm=None) -> torch.Tensor:
        """codes [F,C], actions [F,A], noise [F,S,nd] (CUDA fp32) -> losses [3] = (D_loss, G_loss, pair_div_loss),
        a device tensor owned by the trainer (overwritten by the next step; no host sync here).
        comm (ndivplanning_b200.sharded.Comm): data-parallel step — this rank's F codes of a world*F global batch,
        one NCCL all-reduce of the flat gradient per optimiser step; the losses are those of the global batch."""
        _lib.require_cuda(codes, actions, noise)
        F = codes.shape[0]
        codes, actions, noise = _lib.f32c(codes), _lib.f32c(actions), _lib.f32c(noise)
        if codes.shape != (F, self.C) or actions.shape != (F, self.A) or noise.shape != (F, self.S, self.nd):
            raise ValueError(f"expected codes [F,{self.C}], actions [F,{self.A}], noise [F,{self.S},{self.nd}]")
        with torch.cuda.device(self.device):
            if comm is None:
                _lib.check(_lib.lib().ndp_gan_train_step(self._h, _lib.ptr(codes), _lib.ptr(actions), _lib.ptr(noise), F,
                                                         _lib.ptr(self.losses), _lib.stream_ptr(self.device)))
            else:
                _lib.check(_lib.lib().ndp_gan_train_step_dp(self._h, comm._h, comm.world, _lib.ptr(codes), _lib.ptr(actions),
                                                            _lib.ptr(noise), F, _lib.ptr(self.losses),
                                                            _lib.stream_ptr(self.device)))
        return self.losses

    def capture(self, codes: torch.Tensor, actions: torch.Tensor, noise: torch.Tensor) -> "torch.cuda.CUDAGraph":
        """captures one step reading from (and only from) the given static input tensors; replay() after
        refilling them in place.  The optimiser step counts advance on the device, so replays are real steps."""
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                self.step(codes, actions, noise)
        torch.cuda.current_stream(self.device).wait_stream(side)
        self._graph = g
        return g

    # ---- inspection ------------------------------------------------------------------------------------------
    def grads(self):
        """flat gradients of the last step: (generator of G_loss + factor*div, discriminator of D_loss)"""
        ng, ndd = C.c_size_t(), C.c_size_t()
        _lib.check(_lib.lib().ndp_gan_param_counts(self._h, C.byref(ng), C.byref(ndd)))
        g = torch.empty(ng.value, dtype=torch.float32, device=self.device)
        d = torch.empty(ndd.value, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_grads(self._h, _lib.ptr(g), _lib.ptr(d), _lib.stream_ptr(self.device)))
        return g, d

    def last_actions(self, F: int) -> torch.Tensor:
        a = torch.empty(F * self.S, self.A, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_gan_last_actions(self._h, _lib.ptr(a), F, _lib.stream_ptr(self.device)))
        return a


def flatten_grads(module, layers) -> torch.Tensor:
    """module .grad tensors in the flat [W0|b0|W1|b1|...] layout of ndp_gan_grads"""
    return torch.cat([t.grad.reshape(-1) for n in layers for t in (getattr(module, n).weight, getattr(module, n).bias)])


def unpack_trajectories(images: torch.Tensor, actions: torch.Tensor, seq_length: int):
    """train_gan.py:126-137: images [b,seq,...], actions [b,seq,A] -> (state_cur [b*(seq-1),...],
    state_target [b*(seq-1),...] (the last frame of each trajectory, repeated), actions [b*(seq-1),A])."""
    state_cur, state_target = torch.split(images, split_size_or_sections=[seq_length - 1, 1], dim=1)
    state_cur = state_cur.reshape(-1, *state_cur.shape[2:])
    state_target = torch.repeat_interleave(state_target.squeeze(dim=1), repeats=seq_length - 1, dim=0)
    acts = actions[:, :-1].reshape(-1, actions.shape[-1])
    return state_cur, state_target, acts


@torch.no_grad()
def train_gan_epoch(trainer: "GanTrainer", loader, encoder, seq_length: int, noise_fn=None):
    """One epoch of the reference's GAN training loop (train_gan.py:114-207) with the loop body on the device:
    per batch — unpack the trajectories (:126-137), encode current and target frames with the (frozen, detached)
    `encoder` (:152-155), draw the noise (:43-47; `noise_fn(F) -> [F,S,nd]` overrides the device generator), one
    `trainer.step`.  The three losses are accumulated on the device (the reference syncs `.cpu()` three times per
    batch, :203-205) and returned as (D_loss_avg, G_loss_avg, pair_div_loss_avg) (:207-209), one sync at the end."""
    dev = trainer.device
    sums = torch.zeros(3, dtype=torch.float32, device=dev)
    n_batches = 0
    for images, _states, actions, _goal in loader:
        images, actions = images.float().to(dev), actions.float().to(dev)
        state_cur, state_target, acts = unpack_trajectories(images, actions, seq_length)
        if isinstance(encoder, FusedEncoder):
            # frames -> codes on the device (csrc/k5_encoder.cu); each trajectory's target frame is encoded once and
            # its code repeated, where the reference encodes the seq-1 repeated copies (:133, :153)
            tgt = encoder(images[:, -1]).repeat_interleave(seq_length - 1, dim=0)
            codes = torch.cat([encoder(state_cur), tgt], dim=1).reshape(state_cur.shape[0], -1)
        else:
            codes = torch.cat([encoder(state_cur), encoder(state_target)], dim=1).reshape(state_cur.shape[0], -1)
        F = codes.shape[0]
        noise = noise_fn(F).to(dev) if noise_fn is not None else torch.rand(F, trainer.S, trainer.nd, device=dev)
        sums += trainer.step(codes, acts, noise)
        n_batches += 1
    avg = (sums / max(n_batches, 1)).cpu()
    return float(avg[0]), float(avg[1]), float(avg[2])
