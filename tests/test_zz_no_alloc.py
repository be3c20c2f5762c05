"""hugs_b200_inclusive_scan_static: the prefix sum with its look-back state in a file-scope device array of the
library instead of caller-provided temp storage -- an entry point that allocates nothing (for CUDA-graph
capture, serving loops, harnesses that forbid cudaMalloc; no counterpart in the reference).  Through the C ABI,
against the existing hugs_b200_inclusive_scan: bit-identical (integer work), on calls in a row that share the
array (a larger scan first, so the second call finds stale state of more tiles than it uses).

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _input(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 40, n).astype(np.uint32)
    pad = (-n) % 4
    return x, torch.from_numpy(np.concatenate([x, np.zeros(pad, np.uint32)]).view(np.int32)).cuda()


def _scan_with_temp(N, xin, n):
    out = torch.empty_like(xin)
    tb = int(N.scan_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    N.check(N.inclusive_scan(xin.data_ptr(), out.data_ptr(), n, temp.data_ptr(), tb, _stream()), "inclusive_scan")
    return out


def test_static_scan_equals_the_scan_with_caller_temp_on_calls_in_a_row():
    from hugs_b200 import _native as N
    assert N.inclusive_scan_static_max_items() == 1 << 24
    _, w = _input(4096, 0)   # first use on this device: the library looks the array's address up once
    w_out = torch.empty_like(w)
    N.check(N.inclusive_scan_static(w.data_ptr(), w_out.data_ptr(), 4096, _stream()), "warm-up")
    for n, seed in ((3_000_017, 1), (1_110_210, 2), (8193, 3), (1_110_210, 4), (1, 5), (1 << 24, 6)):
        x, xin = _input(n, seed)
        want = _scan_with_temp(N, xin, n)
        out = torch.full_like(xin, -1)
        torch.cuda.synchronize()
        reserved, free = torch.cuda.memory_reserved(), torch.cuda.mem_get_info()[0]
        N.check(N.inclusive_scan_static(xin.data_ptr(), out.data_ptr(), n, _stream()), "inclusive_scan_static")
        torch.cuda.synchronize()
        # nothing was allocated: neither through torch nor behind its back
        assert torch.cuda.memory_reserved() == reserved and torch.cuda.mem_get_info()[0] == free
        assert torch.equal(out[:n], want[:n]), n
        got = out.cpu().numpy().view(np.uint32)[:n]
        assert (got == np.cumsum(x, dtype=np.uint64).astype(np.uint32)).all(), n


def test_static_scan_back_to_back_on_one_stream_without_a_sync_in_between():
    from hugs_b200 import _native as N
    (xa, a), (xb, b) = _input(2_500_000, 11), _input(700_001, 12)
    oa, ob = torch.empty_like(a), torch.empty_like(b)
    for _ in range(3):   # the same stream orders the calls: the second call's clear follows the first call's kernel
        N.check(N.inclusive_scan_static(a.data_ptr(), oa.data_ptr(), 2_500_000, _stream()), "inclusive_scan_static")
        N.check(N.inclusive_scan_static(b.data_ptr(), ob.data_ptr(), 700_001, _stream()), "inclusive_scan_static")
    torch.cuda.synchronize()
    assert torch.equal(oa[:2_500_000], _scan_with_temp(N, a, 2_500_000)[:2_500_000])
    assert torch.equal(ob[:700_001], _scan_with_temp(N, b, 700_001)[:700_001])


def test_static_scan_rejects_sizes_beyond_the_maximum():
    from hugs_b200 import _native as N
    x = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = N.inclusive_scan_static(x.data_ptr(), x.data_ptr(), (1 << 24) + 1, _stream())   # rejected before any access
    assert rc == -3 and b"exceed the static scratch's maximum" in N.last_error()
    assert N.inclusive_scan_static(None, None, 5, None) == -1
    assert N.inclusive_scan_static(None, None, -1, None) == -1
    assert N.inclusive_scan_static(None, None, 0, None) == 0
    torch.cuda.synchronize()
