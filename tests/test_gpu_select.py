"""GPU: K3 segmented argmin/top-k against the reference's strict-< scan (mpc_eval.py:159-165)
restated in the oracle. Bit-exact indices."""
import pytest
import torch

from oracle import ndiv_oracle as O

pytestmark = pytest.mark.gpu


def _check(cost_cpu, topk, idx_base=0):
    from ndivplanning_b200 import select_plan, unpack_index
    cost = cost_cpu.cuda()
    packed, tk_idx, tk_cost = select_plan(cost, topk, idx_base)
    idx = unpack_index(packed).cpu()
    G, K = cost_cpu.shape
    for g in range(G):
        ref = O.select_first_min(cost_cpu[g])
        valid = bool((cost_cpu[g] < 1e10).any())
        assert int(idx[g]) == (ref + idx_base if valid else 0), (g, int(idx[g]), ref)
        if valid:
            assert (int(packed[g].item()) & 0xFFFFFFFFFFFFFFFF) == O.pack_key(float(cost_cpu[g, ref]), ref + idx_base)
        if topk:
            kk = min(topk, K)
            want = O.select_topk(cost_cpu[g], kk)
            got = tk_idx[g, :kk].cpu().long() - idx_base
            assert got.tolist() == want.tolist(), (g, got.tolist()[:8], want.tolist()[:8])
            gc, wc = tk_cost[g, :kk].cpu(), cost_cpu[g][want]
            assert torch.equal(torch.nan_to_num(gc, nan=-7.0), torch.nan_to_num(wc, nan=-7.0))
            assert (tk_idx[g, kk:] == -1).all()


@pytest.mark.parametrize("G,K,topk", [(1, 1, 0), (1, 5, 1), (3, 31, 4), (2, 1000, 16), (1, 4096, 128),
                                      (7, 5000, 8), (1, 65536, 16), (1, 1 << 20, 32), (64, 4096, 1),
                                      (1, 100000, 0), (5, 3, 8)])
def test_select_random(cuda, G, K, topk):
    torch.manual_seed(G * 1000 + K)
    _check(torch.rand(G, K), topk)


def test_select_ties_nan_and_threshold(cuda):
    torch.manual_seed(0)
    c = torch.randint(0, 50, (4, 70000)).float()          # massive ties: first index must win
    _check(c, 16)
    c = torch.rand(3, 9000)
    c[0, ::3] = float("nan"); c[1, :] = float("nan"); c[2, 100:] = float("inf")
    _check(c, 8)
    c = torch.full((2, 3000), 1e10); c[1, 2999] = 9.99e9   # nothing below 1e10 -> index 0
    _check(c, 4)
    c = torch.randn(2, 20000)                               # negative costs order correctly
    c[0, 77] = -0.0; c[0, 78] = 0.0
    _check(c, 16, idx_base=123456)


def test_select_empty_segments(cuda):
    from ndivplanning_b200 import select_plan
    packed, _, _ = select_plan(torch.empty(2, 0, device="cuda"), 0)
    assert [int(v) & 0xFFFFFFFFFFFFFFFF for v in packed.tolist()] == [0xFFFFFFFF00000000] * 2
    packed, _, _ = select_plan(torch.empty(0, 5, device="cuda"), 0)
    assert packed.numel() == 0
