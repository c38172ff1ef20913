"""Device time of the two ndiv passes (BASELINE config 3 and a wide variant), host latency excluded:
a spin kernel runs first so that every launch is already queued when the start event is reached."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import diversity as div, synthetic  # noqa: E402


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        torch.cuda._sleep(1_000_000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)


def main():
    cfgs = ((1, 4096, 32, 16), (1, 4096, 256, 256), (56, 6, 4, 2), (1, 1024, 32, 16), (8, 4096, 32, 16))
    if len(sys.argv) > 1 and sys.argv[1] == "cfg3":
        cfgs = cfgs[:1]
    for N, k, Dx, Dz in cfgs:
        x, z = synthetic.make_ndiv_inputs(N, k, Dx, Dz)
        x, z = x.cuda(), z.cuda()
        sx, sz = div.ndiv_rowsums(x, z)
        t1 = timed(lambda: div.ndiv_rowsums(x, z, sx=sx, sz=sz))
        t2 = timed(lambda: div.ndiv_loss_grad(x, z, sx, sz, True))
        t3 = timed(lambda: div.ndiv_loss_grad(x, z, sx, sz, False))

        def step():
            xg = x.detach().requires_grad_(True)
            div.compute_pairwise_divergence(xg, z).backward()
        t4 = timed(step)
        pairs = N * k * k
        print(f"N={N} k={k} Dx={Dx} Dz={Dz}: rowsums {t1[0]*1e3:.1f} us | loss+grad {t2[0]*1e3:.1f} us | loss only "
              f"{t3[0]*1e3:.1f} us | autograd fwd+bwd {t4[0]*1e3:.1f} us (avg {t4[1]*1e3:.1f}) -> "
              f"{pairs / (t4[0] * 1e-3):.3e} pairs/s", flush=True)


if __name__ == "__main__":
    main()
