"""Drop-in for the reference's ``diversity`` module (reference diversity.py:8-41), backed by
the sm_100a kernels of libndivplan (csrc/k2_ndiv.cu).  Same names, argument order and
results; autograd works through ``torch.autograd.Function``.  CUDA tensors only.

    import ndivplanning_b200.diversity as div
    pair_div_loss = div.compute_pairwise_divergence(action_hat.view(N, S, -1), noises)   # train_gan.py:193-196
"""
import ctypes as C

import torch

from . import _lib
from ._lib import check, f32c, lib, ptr, require_cuda, stream_ptr

MARGIN = 0.8  # diversity.py:40


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 8), dtype=torch.uint8, device=device)


class _Pairwise(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, normalise):
        require_cuda(v)
        vc = f32c(v)
        N, k, Cdim = vc.shape
        out = torch.empty(N, k, k, dtype=torch.float32, device=v.device)
        with torch.cuda.device(v.device):
            check(lib().ndp_pairwise(ptr(vc), N, k, Cdim, int(normalise), ptr(out), stream_ptr(v.device)))
        ctx.save_for_backward(vc)
        ctx.normalise = int(normalise)
        ctx.in_dtype = v.dtype
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (vc,) = ctx.saved_tensors
        N, k, Cdim = vc.shape
        go = f32c(grad_out)
        gv = torch.empty_like(vc)
        rs = torch.empty(N, k, dtype=torch.float32, device=vc.device) if ctx.normalise else None
        with torch.cuda.device(vc.device):
            check(lib().ndp_pairwise_backward(ptr(vc), ptr(go), N, k, Cdim, ctx.normalise, ptr(rs), ptr(gv),
                                              stream_ptr(vc.device)))
        return gv.to(ctx.in_dtype), None


def compute_pairwise(z):
    """diversity.py:8-9 — z [N,k,C] -> Euclidean distance matrices [N,k,k]."""
    return _Pairwise.apply(z, False)


def compute_pair_distance(z, weight=None):
    """diversity.py:12-19 — distances divided by the detached row sum."""
    if weight is not None:  # unused by every reference call site; composed from the plain kernel
        d = _Pairwise.apply(weight, False) * _Pairwise.apply(z, False)
        return d / torch.sum(d, dim=2)[..., None].detach()
    return _Pairwise.apply(z, True)


def compute_pair_unnormal_distance(z, weight=None):
    """diversity.py:22-29 — un-normalised variant (unused by the reference)."""
    if weight is not None:
        return _Pairwise.apply(weight, False) * _Pairwise.apply(z, False)
    return _Pairwise.apply(z, False)


def ndiv_rowsums(x, z, row0=0, rows=None, sx=None, sz=None):
    """Pass 1 of the loss: row sums of both distance matrices for rows [row0,row0+rows)."""
    N, k, Dx = x.shape
    Dz = z.shape[2]
    rows = k - row0 if rows is None else rows
    if sx is None:
        sx = torch.empty(N, k, dtype=torch.float32, device=x.device)
        sz = torch.empty(N, k, dtype=torch.float32, device=x.device)
    nbytes = C.c_size_t()
    check(lib().ndp_ndiv_workspace_bytes(N, k, Dx, Dz, C.byref(nbytes)))
    ws = _ws(nbytes.value, x.device)
    with torch.cuda.device(x.device):
        check(lib().ndp_ndiv_rowsums(ptr(x), ptr(z), N, k, Dx, Dz, row0, rows, ptr(sx), ptr(sz), ptr(ws), ws.numel(),
                                     stream_ptr(x.device)))
    return sx, sz


def ndiv_loss_grad(x, z, sx, sz, want_grad, row0=0, rows=None, margin=MARGIN):
    """Pass 2: (loss[1], grad_x[N,k,Dx] or None) for rows [row0,row0+rows); grad rows outside stay 0."""
    N, k, Dx = x.shape
    Dz = z.shape[2]
    rows = k - row0 if rows is None else rows
    loss = torch.empty(1, dtype=torch.float32, device=x.device)
    full = (row0 == 0 and rows == k)
    grad = None
    if want_grad:
        grad = torch.empty_like(x) if full else torch.zeros_like(x)
    nbytes = C.c_size_t()
    check(lib().ndp_ndiv_workspace_bytes(N, k, Dx, Dz, C.byref(nbytes)))
    ws = _ws(nbytes.value, x.device)
    with torch.cuda.device(x.device):
        check(lib().ndp_ndiv_loss_grad(ptr(x), ptr(z), ptr(sx), ptr(sz), N, k, Dx, Dz, margin, row0, rows,
                                       ptr(loss), ptr(grad), ptr(ws), ws.numel(), stream_ptr(x.device)))
    return loss, grad


def ndiv_forward_backward(x, z, want_grad, margin=MARGIN):
    """Both passes in one C call (ndp_ndiv_forward_backward): (loss[1], grad_x or None)."""
    N, k, Dx = x.shape
    Dz = z.shape[2]
    dev = x.device
    sx = torch.empty(N, k, dtype=torch.float32, device=dev)
    sz = torch.empty(N, k, dtype=torch.float32, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    grad = torch.empty_like(x) if want_grad else None
    nbytes = C.c_size_t()
    check(lib().ndp_ndiv_workspace_bytes(N, k, Dx, Dz, C.byref(nbytes)))
    ws = _ws(nbytes.value, dev)
    with torch.cuda.device(dev):
        check(lib().ndp_ndiv_forward_backward(ptr(x), ptr(z), N, k, Dx, Dz, margin, ptr(sx), ptr(sz), ptr(loss), ptr(grad),
                                              ptr(ws), ws.numel(), stream_ptr(dev)))
    return loss, grad


class _NdivLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, recodes, codes):
        require_cuda(recodes, codes)
        x, z = f32c(recodes), f32c(codes)
        want_x, want_z = bool(ctx.needs_input_grad[0]), bool(ctx.needs_input_grad[1])
        N, k, Dx = x.shape
        Dz = z.shape[2]
        dev = x.device
        sx = torch.empty(N, k, dtype=torch.float32, device=dev)
        sz = torch.empty(N, k, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        grad = torch.empty_like(x) if want_x else None
        nbytes = C.c_size_t()
        check(lib().ndp_ndiv_workspace_bytes(N, k, Dx, Dz, C.byref(nbytes)))
        ws = _ws(nbytes.value, dev)
        with torch.cuda.device(dev):
            check(lib().ndp_ndiv_forward_backward(ptr(x), ptr(z), N, k, Dx, Dz, MARGIN, ptr(sx), ptr(sz), ptr(loss),
                                                  ptr(grad), ptr(ws), ws.numel(), stream_ptr(dev)))
        ctx.save_for_backward(*([grad] if want_x else []), *([x, z, sx, sz] if want_z else []))
        ctx.want = (want_x, want_z)
        ctx.in_dtypes = (recodes.dtype, codes.dtype)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, grad_out):
        want_x, want_z = ctx.want
        saved = list(ctx.saved_tensors)
        gx = gz = None
        if want_x:
            gx = (saved.pop(0) * grad_out).to(ctx.in_dtypes[0])      # d loss / d recodes was produced with the forward
        if want_z:
            # d loss / d codes (what autograd gives when codes.requires_grad, diversity.py:36-41): exact secondary path
            x, z, sx, sz = saved
            N, k, Dx = x.shape
            Dz = z.shape[2]
            dev = x.device
            nbytes = C.c_size_t()
            check(lib().ndp_ndiv_gradz_workspace_bytes(N, k, Dx, Dz, C.byref(nbytes)))
            ws = _ws(nbytes.value, dev)
            gz = torch.empty_like(z)
            scratch = torch.empty(1, dtype=torch.float32, device=dev)
            gl = f32c(grad_out).reshape(1)
            with torch.cuda.device(dev):
                check(lib().ndp_ndiv_backward(ptr(x), ptr(z), ptr(sx), ptr(sz), N, k, Dx, Dz, MARGIN, ptr(gl), None, ptr(gz),
                                              ptr(scratch), ptr(ws), ws.numel(), stream_ptr(dev)))
            gz = gz.to(ctx.in_dtypes[1])
        return gx, gz


def compute_pairwise_divergence(recodes, codes):
    """diversity.py:36-41 — ``relu(0.8 * delta_z - delta_x).sum()``; outputs first, latents second.  Differentiable
    w.r.t. both arguments like the reference under autograd (the reference's own call site only needs recodes)."""
    N, k, _ = codes.size()
    z = torch.squeeze(codes).view(N, k, -1)
    x = torch.squeeze(recodes).view(N, k, -1)
    return _NdivLoss.apply(x, z)
