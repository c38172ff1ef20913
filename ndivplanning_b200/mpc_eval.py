"""Drop-in for ``mpc_eval.fetch_push_control_evaluation`` (reference mpc_eval.py:39-191) in
embedding mode: same argument order, same duck-typed ``dataset`` / ``config``, same return
tuple ``(avg_action_error, avg_image_loss)``.  The inner loop (mpc_eval.py:124-165) runs on
the fused sm_100a kernels through ``FusedPlanner``; the receding-horizon bookkeeping around
it reproduces the reference's quirks.

Embedding mode = ``image_encoder`` is ``nn.Identity`` and dataset "images" are [seq,E,1,1]
(or [seq,E]) embeddings.  Frame mode (first slice) = ``image_encoder`` is a reference
``image_autoencoder.Encoder`` / ``ImageEncoder`` / ``FusedEncoder`` and dataset images are [seq,3,128,128]
frames: they are encoded once per trajectory by csrc/k5_encoder.cu and the loop plans on the codes with the
embedding-space forward model.  The reference's image-space U-net forward model stays out of scope.
"""
import numpy as np
import torch
from torch import nn

from . import _lib
from .planner import FusedPlanner


def sample_noise(T: int, K: int, nd: int, num_sample: int = 1) -> torch.Tensor:
    """The reference sampler (mpc_eval.py:78-83): one CPU `FloatTensor(K,S,nd).uniform_()` per
    rollout step, in step order.  Returns [T,K,nd] (S=1)."""
    return torch.stack([torch.FloatTensor(K, num_sample, nd).uniform_()[:, 0] for _ in range(T)], 0)


def sample_noise_device(T: int, G: int, K: int, nd: int, seed: int, offset: int = 0, device=None) -> torch.Tensor:
    """The sampler next to its consumer: noise [T,G,K,nd] ~ U[0,1) drawn on the GPU by the library's counter-based
    Philox4x32-10 (ndp_sample_uniform) — reproducible from (seed, offset), no host RNG, no H2D copy.  Use a fresh
    `offset` (or seed) per plan.  Not bit-compatible with the reference's CPU mt19937 stream (nothing on a GPU is):
    parity tests feed the reference's own draws through `sample_noise` instead."""
    import ctypes as C
    dev = torch.device(device if device is not None else torch.cuda.current_device())
    if dev.type != "cuda":
        raise _lib.NdpError("sample_noise_device draws on the GPU; there is no CPU fallback")
    out = torch.empty(T, G, K, nd, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().ndp_sample_uniform(C.c_uint64(seed & (2 ** 64 - 1)), C.c_uint64(offset & (2 ** 64 - 1)),
                                                 _lib.ptr(out), out.numel(), _lib.stream_ptr(dev)))
    return out


def fetch_push_control_evaluation(image_encoder, fwd_model_autoencoder, generator, dataset, config,
                                  return_indices: bool = False):
    if not torch.cuda.is_available():
        raise _lib.NdpError("CUDA is not available; there is no CPU fallback")
    from .encoder import FusedEncoder
    fused_encoder = None
    if isinstance(image_encoder, FusedEncoder):
        fused_encoder = image_encoder
    elif isinstance(image_encoder, nn.Module) and hasattr(image_encoder, "conv6"):
        image_encoder.eval()                                                  # mpc_eval.py:58
        fused_encoder = "build"
    elif not isinstance(image_encoder, nn.Identity):
        raise _lib.NdpError("image_encoder must be nn.Identity() (dataset images are embeddings), an "
                            "image_autoencoder.Encoder-layout module or a FusedEncoder (dataset images are frames)")
    else:
        image_encoder.eval()
    fwd_model_autoencoder.eval(); generator.eval()                            # mpc_eval.py:59-60
    num_sample = config.evaluation.num_sample
    nd = config.evaluation.noise_dim
    K = config.mpc.rollouts
    Th = config.mpc.time_horizon
    if num_sample != 1 or config.evaluation.batch_size != 1:
        raise _lib.NdpError("the reference loop is only well-formed for num_sample=1, batch_size=1 "
                            "(mpc_eval.py:155,161)")
    dev = torch.device("cuda", int(config.gpu_id) if str(config.gpu_id).isdigit() else torch.cuda.current_device())
    if fused_encoder == "build":
        fused_encoder = FusedEncoder(image_encoder, max_images=max(2, int(dataset.seq_length)), device=dev)
    torch.manual_seed(config.random_seed)                                     # mpc_eval.py:75-76
    np.random.seed(config.random_seed)

    planners = {}

    def planner_for(T):
        if T not in planners:
            planners[T] = FusedPlanner(generator, fwd_model_autoencoder, horizon=T, device=dev)
        return planners[T]

    seq = dataset.seq_length
    n_traj = len(dataset)
    action_error_sum = torch.zeros((), device=dev)
    image_error_sum = torch.zeros((), device=dev)
    all_best = []
    for i in range(n_traj):
        images, _states, actions, _goal = dataset[i]
        images = torch.as_tensor(images).float().to(dev)
        if fused_encoder is not None:
            # frames -> codes ONCE per trajectory (the reference re-encodes the K-times repeated goal inside every
            # rollout step, mpc_eval.py:135-136; the codes are the same every time)
            images = fused_encoder(images.reshape(seq, 3, 128, 128))
        images = images.reshape(seq, -1)                                       # [seq,E]
        actions = torch.as_tensor(actions).float().to(dev)[:-1]                # mpc_eval.py:107
        state_cur, goal = images[:-1], images[-1:]                             # mpc_eval.py:103-105
        image_error_sum = torch.zeros((), device=dev)                          # reset per trajectory (:110)
        cur = state_cur[0:1]
        chosen, best_row = [], []
        for image_num in range(seq - 1):
            state_fut = state_cur[image_num + 1:image_num + 2] if image_num != seq - 2 else goal
            T_eff = min(Th, seq - 1 - image_num)                               # mpc_eval.py:127
            noise = sample_noise(T_eff, K, nd).to(dev)[:, None]                # [T,1,K,nd]
            pl = planner_for(T_eff)
            plan = pl.plan(cur, goal, noise)
            best_action = plan.best_action                                     # [1,4]   (:165)
            cur = pl.forward_model(cur, best_action)                           # (:167-169)
            chosen.append(best_action)
            best_row.append(plan.best_idx)
            image_error_sum = image_error_sum + torch.mean((cur - state_fut) ** 2)   # (:173-176)
        action_hat = torch.cat(chosen, dim=0)
        action_error_sum = action_error_sum + torch.mean((actions - action_hat) ** 2)  # (:178-183)
        all_best.append(torch.cat(best_row))
    denom = (seq - 1) * n_traj                                                 # (:185-186)
    out = (float(action_error_sum / denom), float(image_error_sum / denom))
    if return_indices:
        return out + (torch.stack(all_best).cpu(),)
    return out
