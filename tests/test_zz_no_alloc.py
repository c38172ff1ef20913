"""ndp_select_static — the allocation-free variant of ndp_select (scratch in a file-scope static __device__ array; no
counterpart in the reference) — called through the C-ABI and compared
with ndp_select on the same inputs: integer / index work, so the bar is bit-exact.  Two calls in a row on different
inputs (the shared array keeps the last call's contents: the second call must not see the first one's), multi-chunk
shapes (ticket words in the scratch), and the size limit.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

from ndivplanning_b200._lib import NdpError, check, lib, ptr, stream_ptr

pytestmark = pytest.mark.gpu


def _select(cost, topk, idx_base, static):
    G, K = cost.shape
    dev = cost.device
    packed = torch.empty(G, dtype=torch.int64, device=dev)
    ti = torch.full((G, max(topk, 1)), -7, dtype=torch.int32, device=dev)
    tc = torch.full((G, max(topk, 1)), -7.0, dtype=torch.float32, device=dev)
    L = lib()
    with torch.cuda.device(dev):
        if static:
            check(L.ndp_select_static(ptr(cost), G, K, topk, idx_base, ptr(packed), ptr(ti) if topk else None,
                                      ptr(tc) if topk else None, stream_ptr(dev)))
        else:
            n = C.c_size_t()
            check(L.ndp_select_workspace_bytes(G, K, topk, C.byref(n)))
            ws = torch.empty(max(n.value, 256), dtype=torch.uint8, device=dev)
            check(L.ndp_select(ptr(cost), G, K, topk, idx_base, ptr(packed), ptr(ti) if topk else None,
                               ptr(tc) if topk else None, ptr(ws), ws.numel(), stream_ptr(dev)))
    torch.cuda.synchronize(dev)
    return packed, ti, tc


@pytest.mark.parametrize("G,K,topk", [(1, 65536, 16), (3, 20000, 128), (64, 4096, 0), (2, 1 << 20, 16), (5, 77, 8)])
def test_select_static_equals_select_two_calls(cuda, G, K, topk):
    n = C.c_size_t()
    check(lib().ndp_select_static_max_bytes(C.byref(n)))
    assert n.value == 1 << 20
    for call in range(2):                                  # second call: other inputs, same shared scratch
        torch.manual_seed(1000 * call + G + K)
        cost = torch.rand(G, K, device="cuda")
        if call == 1:
            cost[0, K // 2] = cost[0, K - 1] = -1.0        # cross-chunk tie: first index wins
            if G > 1:
                cost[1, :] = float("nan")
        want = _select(cost, topk, 11, static=False)
        got = _select(cost, topk, 11, static=True)
        assert torch.equal(got[0], want[0])
        assert torch.equal(got[1], want[1])
        assert torch.equal(torch.nan_to_num(got[2], nan=-3.0), torch.nan_to_num(want[2], nan=-3.0))
        amin = torch.argmin(torch.nan_to_num(cost[0], nan=float("inf"))).item()
        assert (int(got[0][0].item()) & 0xFFFFFFFF) == amin + 11


def test_select_static_rejects_beyond_its_scratch(cuda):
    """G=4096, K=65,536, top-16 needs 8.4 MB of survivors: refused with a clear error before anything is launched
    (the dummy 4-byte cost buffer is never read)"""
    dummy = torch.zeros(1, device="cuda")
    out = torch.zeros(4096, dtype=torch.int64, device="cuda")
    rc = lib().ndp_select_static(ptr(dummy), 4096, 65536, 16, 0, ptr(out), ptr(out), ptr(out), stream_ptr(dummy.device))
    assert rc != 0
    assert "static array" in lib().ndp_last_error().decode()
    with pytest.raises(NdpError):
        check(rc)
    torch.cuda.synchronize()
