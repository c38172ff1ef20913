"""Gram (tensor-core) ndiv path vs the fp64 oracle and vs forced exact path, plus the ill-conditioned fallback."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import diversity as div, synthetic
from oracle import ndiv_oracle as O

def run(x, z):
    xg = x.cuda().requires_grad_(True)
    loss = div.compute_pairwise_divergence(xg, z.cuda())
    loss.backward()
    return loss.item(), xg.grad.cpu()

print("exact forced" if os.environ.get("NDP_NDIV_EXACT") else "gram fast path")
for (N, k, Dx, Dz, tag) in [(1, 1024, 32, 16, "generic"), (2, 300, 32, 16, "ragged"), (1, 4096, 32, 16, "cfg3"), (1, 512, 4, 2, "narrow")]:
    x, z = synthetic.make_ndiv_inputs(N, k, Dx, Dz, seed=k)
    sx, sz = div.ndiv_rowsums(x.cuda(), z.cuda())
    d = torch.cdist(x.double(), x.double()); sx64 = d.sum(-1)
    dz_ = torch.cdist(z.double(), z.double()); sz64 = dz_.sum(-1)
    print(tag, "rowsum rel err x %.2e z %.2e" % (((sx.cpu().double() - sx64).abs() / sx64).max(), ((sz.cpu().double() - sz64).abs() / sz64).max()))
    l, g = run(x, z)
    l64, g64 = O.ndiv_fwd_bwd(x.double(), z.double())
    print(tag, "loss rel %.2e grad err/max %.2e" % (abs(l - l64.item()) / abs(l64.item()), (g.double() - g64).abs().max() / g64.abs().max()))
# ill-conditioned: duplicate rows -> must fall back to the exact path (finite grads, exact parity)
x, z = synthetic.make_ndiv_inputs(1, 512, 32, 16, seed=3)
x[0, 100] = x[0, 7]; x[0, 300] = x[0, 7] + 1e-4
l, g = run(x, z)
l64, g64 = O.ndiv_fwd_bwd(x.double(), z.double())
print("duplicates: loss rel %.2e grad err/max %.2e finite %s" % (abs(l - l64.item()) / abs(l64.item()), (g.double() - g64).abs().max() / g64.abs().max(), bool(torch.isfinite(g).all())))
