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
                                      (1, 100000, 0), (5, 3, 8), (300, 4096, 0), (1000, 1001, 0), (128, 37, 0), (200, 4096, 16),
                                      (90, 3000, 128)])
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


def test_select_many_goals_special_values(cuda):
    """the persistent argmin kernel of the many-goal case (G >= 64, K <= 4096): NaN rows, nothing below 1e10, ties,
    negative costs, idx_base"""
    torch.manual_seed(3)
    c = torch.rand(130, 4096)
    c[0, :] = float("nan"); c[1, :] = 1e10; c[2, 5] = c[2, 4000] = -3.0; c[3, ::2] = float("nan"); c[4, 100:] = float("inf")
    c[5] = torch.randint(0, 4, (4096,)).float()
    _check(c, 0, idx_base=777)
    _check(c, 16)


def test_select_empty_segments(cuda):
    from ndivplanning_b200 import select_plan
    packed, _, _ = select_plan(torch.empty(2, 0, device="cuda"), 0)
    assert [int(v) & 0xFFFFFFFFFFFFFFFF for v in packed.tolist()] == [0xFFFFFFFF00000000] * 2
    packed, _, _ = select_plan(torch.empty(0, 5, device="cuda"), 0)
    assert packed.numel() == 0


@pytest.mark.parametrize("P,G,K,topk", [(2, 3, 1001, 8), (8, 2, 4096, 16), (4, 5, 64, 0), (8, 1, 40, 128), (3, 4, 7, 16)])
def test_shard_records_merge_equals_global_selection(cuda, P, G, K, topk):
    """P emulated ranks on one GPU: local K3 per shard -> ndp_shard_pack -> stacked records ->
    ndp_shard_merge must equal the reference scan / stable top-k over ALL K candidates
    (mpc_eval.py:159-165), incl. cross-rank ties, an all-NaN goal and short shards (K/P < topk)."""
    from ndivplanning_b200 import select_plan
    from ndivplanning_b200.planner import Plan
    from ndivplanning_b200.sharded import merge_records, pack_records, shard_range
    torch.manual_seed(P * 100 + K)
    cost = torch.rand(G, K)
    if G > 1:
        cost[1, :] = float("nan")
    if G > 2:
        cost[2, 1] = cost[2, K - 2] = -1.0
    a0 = torch.randn(G, K, 4)
    recs = []
    for r in range(P):
        start, count = shard_range(K, r, P)
        loc = cost[:, start:start + count].contiguous().cuda()
        packed, tk_idx, tk_cost = select_plan(loc, topk, start)
        li = ((packed & 0xFFFFFFFF) - start).clamp(0, max(count - 1, 0))
        li = torch.where(packed == -0x100000000, torch.zeros_like(li), li)
        act = a0[:, start:start + count].cuda()[torch.arange(G), li] if count else torch.zeros(G, 4).cuda()
        ver = torch.full((G,), 1 if r != 1 else 0, dtype=torch.int32).cuda()
        recs.append(pack_records(Plan(None, act.contiguous(), packed, loc, None, tk_idx, tk_cost, None, ver), topk))
    packed, best_idx, best_action, tk_idx, tk_cost, verified = merge_records(torch.stack(recs).contiguous(), topk)
    for g in range(G):
        valid = bool((cost[g] < 1e10).any())
        want = O.select_first_min(cost[g]) if valid else 0
        assert int(best_idx[g]) == want
        if valid:
            assert (int(packed[g].item()) & 0xFFFFFFFFFFFFFFFF) == O.pack_key(float(cost[g, want]), want)
        assert torch.equal(best_action[g].cpu(), a0[g, want])
        if topk:
            kk = min(topk, K)
            wt = O.select_topk(cost[g], kk)
            assert tk_idx[g, :kk].cpu().tolist() == wt.tolist()
            assert torch.equal(torch.nan_to_num(tk_cost[g, :kk].cpu(), nan=-7.0), torch.nan_to_num(cost[g][wt], nan=-7.0))
            assert (tk_idx[g, kk:] == -1).all()
    assert verified.cpu().tolist() == [0 if P > 1 else 1] * G


@pytest.mark.parametrize("G,K,topk", [(3, 1001, 8), (1, 65536, 16), (5, 64, 0), (2, 300, 128)])
def test_shard_exchange_entry_world1(cuda, G, K, topk):
    """ndp_shard_exchange on a world-size-1 communicator runs the peer-window kernels (pack+push into the window,
    wait+merge from it) exactly as an N-rank call does: the global plan must equal the local one, over several
    back-to-back exchanges (slot reuse after 8)"""
    from ndivplanning_b200 import select_plan
    from ndivplanning_b200.planner import Plan
    from ndivplanning_b200.sharded import Comm, exchange_plan, merge_records, pack_records
    comm = Comm.solo("cuda:0")
    try:
        for it in range(11):
            torch.manual_seed(G * 1000 + K + it)
            cost = torch.rand(G, K)
            if G > 1 and it % 2:
                cost[1, :] = float("nan")
            loc = cost.cuda()
            packed, tk_idx, tk_cost = select_plan(loc, topk, 0)
            act = torch.randn(G, 4).cuda()
            ver = torch.randint(0, 2, (G,), dtype=torch.int32).cuda()
            p = Plan(None, act, packed, loc, None, tk_idx, tk_cost, None, ver)
            got = exchange_plan(p, topk, comm)
            want = merge_records(pack_records(p, topk).unsqueeze(0).contiguous(), topk)
            for a, b in zip(got, want):
                if a is None:
                    assert b is None
                else:
                    assert torch.equal(torch.nan_to_num(a.float(), nan=-7.0) if a.is_floating_point() else a,
                                       torch.nan_to_num(b.float(), nan=-7.0) if b.is_floating_point() else b)
        assert comm.peer_status() is True
    finally:
        comm.close()
