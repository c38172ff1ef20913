"""Device time of K3 (ndp_select) alone: BASELINE cfg2 / cfg4-shard (one goal, many candidates) and cfg5 (4096 goals x
4096 candidates = 67 MB, the HBM-bound case), argmin only and with top-16 / top-128, next to torch.argmin / torch.topk."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import select_plan  # noqa: E402


def timed(fn, reps=20):
    fn(); fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(1_000_000)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    shapes = ((1, 65536), (1, 131072), (1, 1 << 20), (64, 4096), (4096, 4096))
    if os.environ.get("NDP_SELECT_ONLY"):                  # e.g. 4096x4096 (profiling runs)
        shapes = (tuple(int(v) for v in os.environ["NDP_SELECT_ONLY"].split("x")),)
    for G, K in shapes:
        bufs = [torch.rand(G, K, device="cuda") for _ in range(3 if G * K * 4 > (32 << 20) else 1)]
        it = [0]

        def nxt():
            it[0] += 1
            return bufs[it[0] % len(bufs)]
        line = f"G={G} K={K}:"
        for tk in (0, 16, 128):
            t = timed(lambda: select_plan(nxt(), tk))
            line += f"  top{tk} {t * 1e3:7.1f} us ({G * K * 4 / (t * 1e-3) / 1e9:6.0f} GB/s)"
        ta = timed(lambda: torch.argmin(nxt(), dim=1))
        tt = timed(lambda: torch.topk(nxt(), 16, dim=1, largest=False))
        line += f"  | torch argmin {ta * 1e3:.1f} us, topk16 {tt * 1e3:.1f} us"
        print(line, flush=True)
        # correctness against torch (stable sort order)
        c = bufs[0]
        pm, ti, tc = select_plan(c, 16)
        want = torch.sort(c, dim=1, stable=True).indices[:, :16]
        assert torch.equal(ti.long(), want), "top-16 mismatch"
        assert torch.equal((pm & 0xFFFFFFFF), torch.argmin(c, dim=1)), "argmin mismatch"


if __name__ == "__main__":
    main()
