"""pinq_compact_flags_static -- the allocation-free variant of pinq_compact_flags (scratch in a static device array inside the
library; no counterpart in the reference) -- called through the C-ABI and checked
against the existing entry point: survivors' indices and count bit-exact (integer work), on two calls in a row, the second
on other data and a smaller n so that stale contents of the shared array would show.  Collected last on purpose (file name):
a failure here must not hide other tests under `pytest -x`.  Works alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _compact_reference(lib, P, flags):
    n = flags.numel()
    out = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
    cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    scratch = torch.empty(int(lib.pinq_compact_scratch_bytes(n)), dtype=torch.uint8, device="cuda")
    P.check(lib.pinq_compact_flags(P.ptr(flags), n, P.ptr(out), P.ptr(cnt), P.ptr(scratch), P.stream()))
    return out, cnt


def test_static_scratch_compaction_matches_the_caller_scratch_entry_point():
    from pin_slam_b200 import _pinq as P
    lib = P.load()
    g = torch.Generator().manual_seed(21)
    # two calls in a row: a ragged size just below a tile multiple with ~37 % set flags, then a smaller, denser batch
    for n, p_set in ((3 * 1024 * 1024 - 5, 0.37), (70_001, 0.9)):
        flags = (torch.rand(n, generator=g) < p_set).to(torch.uint8).cuda()
        want_idx, want_cnt = _compact_reference(lib, P, flags)
        # outputs are allocated BEFORE the call under test; the call itself gets only inputs, outputs and the stream
        out = torch.full((n,), -7, dtype=torch.int32, device="cuda")
        cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        P.check(lib.pinq_compact_flags_static(P.ptr(flags), n, P.ptr(out), P.ptr(cnt), P.stream()))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before             # nothing allocated through torch either
        k = int(want_cnt.item())
        assert int(cnt.item()) == k == int(flags.sum().item())
        assert torch.equal(out[:k], want_idx[:k])                   # bit-exact, order-preserving
        assert torch.equal(out[:k].long(), torch.nonzero(flags).squeeze(1))
        assert bool((out[k:] == -7).all())                          # nothing written past the survivors


def test_static_scratch_limits_and_edge_cases():
    from pin_slam_b200 import _pinq as P
    lib = P.load()
    cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    one = torch.ones(1, dtype=torch.uint8, device="cuda")
    out = torch.empty(1, dtype=torch.int32, device="cuda")
    # beyond the stated maximum: rejected with a clear error before anything is launched or dereferenced
    rc = lib.pinq_compact_flags_static(P.ptr(one), P.COMPACT_STATIC_MAX_N + 1, P.ptr(out), P.ptr(cnt), P.stream())
    assert rc != 0
    msg = lib.pinq_last_error().decode()
    assert "PINQ_COMPACT_STATIC_MAX_N" in msg and str(P.COMPACT_STATIC_MAX_N) in msg
    torch.cuda.synchronize()
    assert int(cnt.item()) == -1
    # empty input: count 0
    P.check(lib.pinq_compact_flags_static(None, 0, None, P.ptr(cnt), P.stream()))
    assert int(cnt.item()) == 0
    # exactly the maximum: all-zero flags except the last one
    n = P.COMPACT_STATIC_MAX_N
    flags = torch.zeros(n, dtype=torch.uint8, device="cuda")
    flags[-1] = 1
    flags[12345] = 1
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    P.check(lib.pinq_compact_flags_static(P.ptr(flags), n, P.ptr(out), P.ptr(cnt), P.stream()))
    assert int(cnt.item()) == 2 and out[:2].tolist() == [12345, n - 1]
