"""TEST INFRASTRUCTURE ONLY — not imported by the product package.

CPU restatement (torch on CPU, fp32 with an fp64 twin) of the hot path of
goodmattg/ndivplanning.  Every function cites the reference lines it follows
(paths relative to /root/reference).  Pinned by tests/golden/*.npz which were
produced by running the UNMODIFIED reference (oracle/make_golden.py).

The arithmetic itself lives in PyTorch (nn.Linear/addmm, torch.norm, MSELoss);
the reference pins no torch version so the effective pin is the
torch installed in this image (2.11.0).
"""
from __future__ import annotations

import torch
from torch import nn
from torch.nn import functional as F

__all__ = [
    "compute_pairwise", "compute_pair_distance", "compute_pairwise_divergence",
    "ndiv_grad_analytic", "ndiv_fwd_bwd_chunked", "GeneratorMLP", "LatentForward", "draw_noise",
    "diverse_sampling", "mpc_rollout", "select_first_min", "select_topk",
    "mpc_plan", "pack_key", "closed_loop_eval", "ndiv_fwd_bwd", "ActionDiscriminator", "gan_optimizers",
    "gan_train_step", "gan_train_epoch", "philox4x32_10", "sample_uniform",
]

MARGIN = 0.8          # diversity.py:40
MIN_ERROR_INIT = 1e10  # mpc_eval.py:122


# ----------------------------------------------------------------------------
# ndiv loss — diversity.py:8-41
# ----------------------------------------------------------------------------
def compute_pairwise(z: torch.Tensor) -> torch.Tensor:
    """diversity.py:8-9 — [N,k,C] -> [N,k,k] Euclidean distances (diag exactly 0)."""
    return torch.norm(z[:, :, None, :] - z[:, None, :, :], p=2, dim=3)


def compute_pair_distance(z: torch.Tensor, weight=None) -> torch.Tensor:
    """diversity.py:12-19 — row-normalise by the DETACHED row sum."""
    if weight is not None:
        d = compute_pairwise(weight) * compute_pairwise(z)
    else:
        d = compute_pairwise(z)
    s = torch.sum(d, dim=2)
    return d / s[..., None].detach()


def compute_pairwise_divergence(recodes: torch.Tensor, codes: torch.Tensor) -> torch.Tensor:
    """diversity.py:36-41 — outputs first, latents second; plain sum."""
    N, k, _ = codes.size()
    z_delta = compute_pair_distance(torch.squeeze(codes).view(N, k, -1))
    x_delta = compute_pair_distance(torch.squeeze(recodes).view(N, k, -1))
    return F.relu(z_delta * MARGIN - x_delta).sum()


def ndiv_fwd_bwd(x: torch.Tensor, z: torch.Tensor):
    """Loss + d loss / d x through autograd, exactly as train_gan.py:193-202 uses it."""
    x = x.detach().clone().requires_grad_(True)
    loss = compute_pairwise_divergence(x, z)
    loss.backward()
    return loss.detach(), x.grad.detach()


def ndiv_grad_analytic(x: torch.Tensor, z: torch.Tensor) -> torch.Tensor:
    """Closed-form gradient, used to cross-check autograd in fp64.

    M_ij = 1[0.8 dz_ij/sz_i - dx_ij/sx_i > 0];  W_ij = -(M_ij/sx_i + M_ji/sx_j)
    dL/dx_i = sum_j (W_ij / dx_ij) (x_i - x_j), terms with dx_ij == 0 dropped.
    """
    N, k = x.shape[:2]
    x = x.reshape(N, k, -1)
    z = z.reshape(N, k, -1)
    dx, dz = compute_pairwise(x), compute_pairwise(z)
    sx, sz = dx.sum(2, keepdim=True), dz.sum(2, keepdim=True)
    M = ((MARGIN * dz / sz - dx / sx) > 0).to(x.dtype)
    W = -(M / sx + M.transpose(1, 2) / sx.transpose(1, 2))
    What = torch.where(dx > 0, W / dx, torch.zeros_like(W))
    return x * What.sum(2, keepdim=True) - What @ x


@torch.no_grad()
def ndiv_fwd_bwd_chunked(x: torch.Tensor, z: torch.Tensor, block: int = 512, want_grad_z: bool = False):
    """compute_pairwise_divergence + its gradient (diversity.py:36-41 under autograd) for ONE group at sizes where
    the reference's [k,k,C] difference tensor does not fit in RAM (BASELINE cfg3: k=4096 -> 4.3 GB in fp64):
    same formulas as ndiv_grad_analytic, evaluated in row blocks of `block` rows.  x [k,Dx], z [k,Dz] (any float
    dtype; use fp64 for the arbiter).  Returns (loss, grad_x, sx, sz[, grad_z]); pinned against autograd by
    tests/test_oracle_golden.py::test_chunked_ndiv_equals_autograd."""
    k = x.shape[0]
    sx = torch.empty(k, dtype=x.dtype)
    sz = torch.empty(k, dtype=x.dtype)
    for r0 in range(0, k, block):
        r1 = min(k, r0 + block)
        sx[r0:r1] = torch.norm(x[r0:r1, None, :] - x[None, :, :], p=2, dim=2).sum(1)
        sz[r0:r1] = torch.norm(z[r0:r1, None, :] - z[None, :, :], p=2, dim=2).sum(1)
    loss = torch.zeros((), dtype=x.dtype)
    gx = torch.zeros_like(x)
    gz = torch.zeros_like(z)
    for r0 in range(0, k, block):
        r1 = min(k, r0 + block)
        dx = torch.norm(x[r0:r1, None, :] - x[None, :, :], p=2, dim=2)          # [b,k]
        dz = torch.norm(z[r0:r1, None, :] - z[None, :, :], p=2, dim=2)
        h_ij = MARGIN * dz / sz[r0:r1, None] - dx / sx[r0:r1, None]               # row i normalisers
        h_ji = MARGIN * dz / sz[None, :] - dx / sx[None, :]                       # d is symmetric: h_ji uses row j's
        loss = loss + F.relu(h_ij).sum()
        M_ij, M_ji = (h_ij > 0).to(x.dtype), (h_ji > 0).to(x.dtype)
        W = -(M_ij / sx[r0:r1, None] + M_ji / sx[None, :])
        What = torch.where(dx > 0, W / dx, torch.zeros_like(W))
        gx[r0:r1] = x[r0:r1] * What.sum(1, keepdim=True) - What @ x
        if want_grad_z:
            Wz = MARGIN * (M_ij / sz[r0:r1, None] + M_ji / sz[None, :])
            Wzh = torch.where(dz > 0, Wz / dz, torch.zeros_like(Wz))
            gz[r0:r1] = z[r0:r1] * Wzh.sum(1, keepdim=True) - Wzh @ z
    if want_grad_z:
        return loss, gx, sx, sz, gz
    return loss, gx, sx, sz


# ----------------------------------------------------------------------------
# modules
# ----------------------------------------------------------------------------
class GeneratorMLP(nn.Module):
    """models/gan.py:61-86 (`Decoder`): (256+nd)->128->64->128->256->4, ReLU, linear head."""

    def __init__(self, noise_dim: int, cond_dim: int = 256):
        super().__init__()
        self.fc1 = nn.Linear(cond_dim + noise_dim, 128)
        self.fc2 = nn.Linear(128, 64)
        self.fc3 = nn.Linear(64, 128)
        self.fc4 = nn.Linear(128, 256)
        self.fc5 = nn.Linear(256, 4)

    def forward(self, z):
        a = F.relu(self.fc1(z))
        a = F.relu(self.fc2(a))
        a = F.relu(self.fc3(a))
        a = F.relu(self.fc4(a))
        return self.fc5(a)


class LatentForward(nn.Module):
    """Embedding-space forward model with ForwardAutoencoder's call contract
    (models/forward_encoder.py:105-114): cat([code, actions],1) — state first,
    action last; training mode returns the residual, eval mode state + residual.
    Body defined by SURVEY.md §8a-F: fc1 Linear(E+4,H) -> ReLU -> fc2 Linear(H,E).
    """

    def __init__(self, E: int = 128, H: int = 1024, A: int = 4):
        super().__init__()
        self.fc1 = nn.Linear(E + A, H)
        self.fc2 = nn.Linear(H, E)

    def forward(self, state_cur, actions):
        code = state_cur.flatten(1)
        r = self.fc2(F.relu(self.fc1(torch.cat([code, actions], dim=1))))
        r = r.view_as(state_cur)
        if self.training:
            return r
        return state_cur + r


# ----------------------------------------------------------------------------
# sampling — mpc_eval.py:78-83
# ----------------------------------------------------------------------------
def draw_noise(T: int, K: int, nd: int, S: int = 1) -> torch.Tensor:
    """One `torch.FloatTensor(K,S,nd).uniform_()` per rollout step, step-major
    (mpc_eval.py:80 called from :141 inside `for ts`). Uses the global CPU RNG."""
    return torch.stack([torch.FloatTensor(K, S, nd).uniform_() for _ in range(T)], 0)


def diverse_sampling(code: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """mpc_eval.py:78-83 with the noise passed in: noise goes in the LAST columns."""
    S = noise.shape[1]
    code = code[:, None, :].expand(-1, S, -1)
    return torch.cat([code, noise], dim=2)


# ----------------------------------------------------------------------------
# MPC inner loop — mpc_eval.py:119-165
# ----------------------------------------------------------------------------
@torch.no_grad()
def mpc_rollout(gen: nn.Module, fwd: nn.Module, e0: torch.Tensor, goal: torch.Tensor,
                noise: torch.Tensor):
    """e0, goal: [E]; noise: [T,K,nd] -> (state_T [K,E,1,1], a0 [K,1,4]).

    Follows mpc_eval.py:124-157 with image_encoder = nn.Identity():
      :125 repeat state K times; :133 repeat target; :137-139 cat([state,target],1).squeeze();
      :141 diverse_sampling; :146-149 generator(view(-1,C+nd)).view(K,-1,4);
      :150-151 keep step-0 actions; :154-157 state <- fwd(state, a.squeeze(1)).
    """
    T, K, nd = noise.shape
    E = e0.numel()
    state = e0.view(1, E, 1, 1).repeat(K, 1, 1, 1)
    target = goal.view(1, E, 1, 1).repeat(K, 1, 1, 1)
    a0 = None
    for ts in range(T):
        now_codes = torch.cat([state, target], dim=1).squeeze()
        if now_codes.dim() == 1:  # K == 1: the reference's squeeze() would drop the batch dim
            now_codes = now_codes[None]
        dz = diverse_sampling(now_codes, noise[ts].view(K, 1, nd))
        a = gen(dz.view(-1, dz.size(2))).view(K, -1, 4)
        if ts == 0:
            a0 = a
        state = fwd(state, a.squeeze(1))
    return state, a0


@torch.no_grad()
def candidate_costs(state_T: torch.Tensor, goal: torch.Tensor) -> torch.Tensor:
    """mpc_eval.py:161 — nn.MSELoss()(state_T[ro], target[0]) for every ro: mean over E."""
    K = state_T.shape[0]
    return ((state_T.reshape(K, -1) - goal.reshape(1, -1)) ** 2).mean(dim=1)


def select_first_min(costs) -> int:
    """mpc_eval.py:122,159-164 — strict '<' scan from min_error = 1e10, index init 0.
    First minimal index wins; NaN never wins; all >= 1e10 -> 0."""
    min_error = MIN_ERROR_INIT
    best = 0
    for ro, c in enumerate(costs.tolist() if hasattr(costs, "tolist") else costs):
        if c < min_error:
            min_error = c
            best = ro
    return best


def select_topk(costs: torch.Tensor, k: int) -> torch.Tensor:
    """north_star's top-k extension: k smallest under the total
    order (cost, index) == stable ascending sort (NaN sorts last)."""
    return torch.sort(costs, stable=True).indices[:k]


def pack_key(cost: float, idx: int) -> int:
    """(ordered_bits(cost) << 32) | idx — the u64 the multi-GPU min-allreduce uses.
    Ordered bits: flip sign bit for non-negatives, all bits for negatives; NaN -> 0xFFFFFFFF."""
    import struct
    if cost != cost:
        key = 0xFFFFFFFF
    else:
        b = struct.unpack("<I", struct.pack("<f", cost))[0]
        key = (b ^ 0xFFFFFFFF) if (b >> 31) else (b | 0x80000000)
    return (key << 32) | idx


@torch.no_grad()
def mpc_plan(gen, fwd, e0, goal, noise, topk: int = 1):
    """One plan: rollout + terminal cost + first-min + top-k.
    Returns dict(cost[K], best_idx, best_action[4], topk_idx[topk], a0[K,4])."""
    state_T, a0 = mpc_rollout(gen, fwd, e0, goal, noise)
    cost = candidate_costs(state_T, goal)
    best = select_first_min(cost)
    return dict(cost=cost, best_idx=best, best_action=a0[best, 0].clone(),
                topk_idx=select_topk(cost, topk), a0=a0[:, 0].clone(), state_T=state_T.flatten(1))


# ----------------------------------------------------------------------------
# whole evaluation function — mpc_eval.py:94-191 (embedding mode, batch_size=1, S=1)
# ----------------------------------------------------------------------------
@torch.no_grad()
def closed_loop_eval(gen, fwd, images: torch.Tensor, actions: torch.Tensor, K: int, Th: int,
                     nd: int, seed: int, vector_argmin: bool = False):
    """images [n_traj, seq, E], actions [n_traj, seq, 4].  Restates the receding-horizon
    loop (mpc_eval.py:112-176) including its quirks (Appendix B): truncated horizon near the
    end (:127), image_error_sum reset per trajectory (:110) but divided by all (:186).
    Returns (avg_action_error, avg_image_loss, best_indices[n_traj, seq-1])."""
    import numpy as np
    torch.manual_seed(seed)
    np.random.seed(seed)
    mse = nn.MSELoss()
    n_traj, seq, E = images.shape
    action_error_sum = 0
    image_error_sum = 0
    best_all = []
    for i in range(n_traj):
        state_cur = images[i, :-1]
        goal = images[i, -1]
        acts = actions[i, :-1]
        image_error_sum = 0
        chosen = []
        best_row = []
        cur = state_cur[0]
        for image_num in range(seq - 1):
            state_fut = state_cur[image_num + 1] if image_num != seq - 2 else goal
            T_eff = min(Th, seq - 1 - image_num)
            noise = draw_noise(T_eff, K, nd)[:, :, 0, :]
            state_T, a0 = mpc_rollout(gen, fwd, cur, goal, noise)
            cost = candidate_costs(state_T, goal)
            if vector_argmin:
                best = int(torch.argmin(torch.where(cost < MIN_ERROR_INIT, cost,
                                                    torch.full_like(cost, float("inf")))))
            else:
                best = select_first_min(cost)
            best_row.append(best)
            best_action = a0[best]  # [1,4]
            cur = fwd(cur.view(1, E, 1, 1), best_action).view(E)
            chosen.append(best_action)
            image_error_sum = image_error_sum + mse(cur, state_fut)
        action_hat = torch.cat(chosen, dim=0)
        action_error_sum = action_error_sum + mse(acts[None], action_hat)
        best_all.append(best_row)
    denom = (seq - 1) * n_traj
    return float(action_error_sum / denom), float(image_error_sum / denom), best_all


# ----------------------------------------------------------------------------
# GAN training step around the ndiv loss — train_gan.py:126-203
# ----------------------------------------------------------------------------
class ActionDiscriminator(nn.Module):
    """models/gan.py:89-110 (Discriminator), same parameter names."""

    def __init__(self, cond_dim: int = 256, A: int = 4):
        super().__init__()
        self.fc1 = nn.Linear(cond_dim + A, 64)
        self.fc2 = nn.Linear(64, 128)
        self.fc3 = nn.Linear(128, 256)
        self.fc4 = nn.Linear(256, 1)

    def forward(self, action, state_code):
        state_action = torch.cat([action, state_code], dim=1)          # gan.py:102
        h = F.leaky_relu(self.fc1(state_action))
        h = F.leaky_relu(self.fc2(h))
        h = F.leaky_relu(self.fc3(h))
        return self.fc4(h)


def gan_optimizers(decoder, discriminator, lr: float = 2e-4):
    """train_gan.py:98-104 (the encoder group of G_optimizer never gets a gradient: :152-153 detach)."""
    G_opt = torch.optim.Adam([{"params": decoder.parameters()}], lr=lr, betas=(0.5, 0.999))
    D_opt = torch.optim.Adam(discriminator.parameters(), lr=lr, betas=(0.5, 0.999))
    return G_opt, D_opt


def gan_train_step(decoder, discriminator, G_opt, D_opt, codes, actions, noise, div_factor: float = 0.1,
                   discrim_steps_per_gen: int = 1, divergence=compute_pairwise_divergence, probe=None):
    """One iteration of train_gan.py:126-203 on encoded inputs.  codes [F,256] = cat([encoder(state_cur),
    encoder(state_target)], 1) (:152-155); actions [F,4] (:137); noise [F,S,nd] = the uniform draw of
    diverse_sampling (:43-47).  Returns (D_loss, G_loss, pair_div_loss, action_hat) as detached tensors."""
    Fb, S, nd = noise.shape
    bce = nn.BCEWithLogitsLoss()
    action_unsqueeze = torch.repeat_interleave(actions, repeats=S, dim=0)         # :140-142
    codes_unsqueeze = torch.repeat_interleave(codes, repeats=S, dim=0)            # :156
    code = codes[:, None, :].expand(-1, S, -1)                                    # :45
    diverse_codes = torch.cat([code, noise], dim=2)                               # :46
    action_hat = decoder(diverse_codes.reshape(-1, diverse_codes.size(2)))        # :165
    DIV = Fb * S
    ones, zeros = torch.ones(DIV, dtype=codes.dtype), torch.zeros(DIV, dtype=codes.dtype)
    D_loss = torch.zeros(())
    for _ in range(discrim_steps_per_gen):                                        # :172-184
        D_loss = bce(torch.squeeze(discriminator(action_unsqueeze, codes_unsqueeze)), ones) + \
            bce(torch.squeeze(discriminator(action_hat, codes_unsqueeze)), zeros)
        D_opt.zero_grad()
        D_loss.backward(retain_graph=True)
        if probe is not None:   # discriminator gradient of the D step (the G backward below accumulates onto .grad)
            probe["d_grad"] = torch.cat([p.grad.detach().reshape(-1).clone() for p in discriminator.parameters()])
        D_opt.step()
    G_loss = bce(torch.squeeze(discriminator(action_hat, codes_unsqueeze)), ones)  # :187-190
    pair_div_loss = divergence(action_hat.view(Fb, S, -1), noise)                 # :193-196
    total_loss = G_loss + div_factor * pair_div_loss                              # :198-200
    G_opt.zero_grad()
    total_loss.backward()
    G_opt.step()                                                                  # :201-203
    return D_loss.detach(), G_loss.detach(), pair_div_loss.detach(), action_hat.detach()


# ----------------------------------------------------------------------------
# device sampler — counter-based Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as
# 1, 2, 3", SC'11; Random123 reference constants), as ndp_sample_uniform defines its stream
# ----------------------------------------------------------------------------
def philox4x32_10(counter, key):
    """counter [n,4] uint32, key [2] -> [n,4] uint32 (numpy)."""
    import numpy as np
    c = np.array(counter, dtype=np.uint64).reshape(-1, 4).copy()
    k0, k1 = np.uint64(key[0]), np.uint64(key[1])
    M0, M1, W0, W1, MASK = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0x9E3779B9), np.uint64(0xBB67AE85), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c[:, 0], M1 * c[:, 2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c = np.stack([hi1 ^ c[:, 1] ^ k0, lo1, hi0 ^ c[:, 3] ^ k1, lo0], axis=1)
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c.astype(np.uint32)


def sample_uniform(seed: int, offset: int, n: int):
    """ndp_sample_uniform(seed, offset, out, n): value i = word (i & 3) of philox({i/4, offset}, seed), top 24 bits."""
    import numpy as np
    q = np.arange((n + 3) // 4, dtype=np.uint64)
    ctr = np.stack([q & np.uint64(0xFFFFFFFF), q >> np.uint64(32),
                    np.full_like(q, offset & 0xFFFFFFFF), np.full_like(q, (offset >> 32) & 0xFFFFFFFF)], axis=1)
    w = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)).reshape(-1)[:n]
    return (w >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def gan_train_epoch(decoder, discriminator, G_opt, D_opt, loader, encoder, seq_length: int, num_sample: int,
                    noise_dim: int, div_factor: float = 0.1, discrim_steps_per_gen: int = 1, noises=None):
    """train_gan.py:114-209 for one epoch on the CPU (encoder frozen/detached, :152-153).  `noises`: optional list
    of [F,S,nd] tensors, one per batch (else the reference's draw, :43-44).  Returns the three loss averages."""
    D_sum = G_sum = P_sum = 0.0
    for i, (images, states, actions, goal) in enumerate(loader):
        images, actions = images.float(), actions.float()
        state_cur, state_target = torch.split(images, split_size_or_sections=[seq_length - 1, 1], dim=1)      # :127
        state_cur = state_cur.reshape(-1, *(state_cur.size()[2:]))                                             # :131
        state_target = torch.repeat_interleave(state_target.squeeze(dim=1), repeats=seq_length - 1, dim=0)     # :133
        acts = actions[:, :-1].reshape(-1, actions.size()[-1])                                                 # :137
        state_codes = encoder(state_cur).detach()                                                              # :152
        target_codes = encoder(state_target).detach()
        codes = torch.cat([state_codes, target_codes], dim=1).squeeze()                                        # :155
        codes = codes.reshape(state_cur.shape[0], -1)
        noise = noises[i] if noises is not None else torch.FloatTensor(codes.shape[0], num_sample, noise_dim).uniform_()
        dl, gl, pl, _ = gan_train_step(decoder, discriminator, G_opt, D_opt, codes, acts, noise, div_factor,
                                       discrim_steps_per_gen)
        D_sum, G_sum, P_sum = D_sum + float(dl), G_sum + float(gl), P_sum + float(pl)
    n = max(len(loader), 1)
    return D_sum / n, G_sum / n, P_sum / n


# ---- image encoder: models/image_autoencoder.py:14-49, eval mode ----------------------
def _conv2d(x, W, b, stride: int, pad: int):
    """nn.Conv2d as im2col + matmul: out[b,o,i,j] = bias[o] + sum_{c,u,v} W[o,c,u,v] x[b,c,i*s-p+u,j*s-p+v]."""
    cols = F.unfold(x, kernel_size=W.shape[2:], stride=stride, padding=pad)          # [B, C*KH*KW, OH*OW]
    OH = (x.shape[2] + 2 * pad - W.shape[2]) // stride + 1
    y = W.flatten(1) @ cols + b.view(1, -1, 1)
    return y.view(x.shape[0], W.shape[0], OH, -1)


def _bn_eval(x, sd, name: str, eps: float = 1e-5):
    """nn.BatchNorm2d in eval mode: (x - running_mean) / sqrt(running_var + eps) * weight + bias."""
    v = lambda k: sd[f"{name}.{k}"].view(1, -1, 1, 1).to(x.dtype)
    return (x - v("running_mean")) / torch.sqrt(v("running_var") + eps) * v("weight") + v("bias")


def encoder_forward(sd, images: torch.Tensor) -> torch.Tensor:
    """Encoder.forward (image_autoencoder.py:35-49) from a state dict: BatchNorm after conv1..conv3 only; conv4_bn and
    conv5_bn are parameters the forward never applies.  Computes in images.dtype."""
    dt = images.dtype
    x = images
    for i in (1, 2, 3, 4, 5):
        x = _conv2d(x, sd[f"conv{i}.weight"].to(dt), sd[f"conv{i}.bias"].to(dt), 2, 1)
        if i <= 3:
            x = _bn_eval(x, sd, f"conv{i}_bn")
        x = F.relu(x)
    return _conv2d(x, sd["conv6.weight"].to(dt), sd["conv6.bias"].to(dt), 1, 0)
