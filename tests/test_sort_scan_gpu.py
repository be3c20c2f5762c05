"""The hand-written onesweep radix sort and decoupled-look-back scan, through the C ABI, against
numpy / torch (CUB's role in the reference: cuda_rasterizer/rasterizer_impl.cu:277,303-308)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sort(keys_np, vals_np, end_bit):
    from hugs_b200 import _native as N

    n = keys_np.shape[0]
    k_in = torch.from_numpy(keys_np.view(np.int64)).cuda()
    v_in = torch.from_numpy(vals_np.view(np.int32)).cuda()
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    tb = int(N.sort_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    rc = N.sort_pairs(N.ptr(k_in), N.ptr(v_in), N.ptr(k_out), N.ptr(v_out), n, end_bit, temp.data_ptr(), tb,
                      C.c_void_p(torch.cuda.current_stream().cuda_stream))
    N.check(rc, "sort_pairs")
    torch.cuda.synchronize()
    return k_out.cpu().numpy().view(np.uint64), v_out.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("n", [0, 1, 2, 31, 4095, 4096, 4097, 8193, 100_003, 3_000_001])
@pytest.mark.parametrize("end_bit", [45, 41, 33])
def test_sort_pairs_stable_bit_exact(n, end_bit):
    rng = np.random.default_rng(n + end_bit)
    tiles = 1 << (end_bit - 32)
    # few distinct depths => many ties: stability is what makes point_list reproducible
    keys = (rng.integers(0, tiles, n).astype(np.uint64) << np.uint64(32)) | \
        rng.integers(0, 1 << 12, n).astype(np.uint64) << np.uint64(18)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = _sort(keys, vals, end_bit)
    order = np.argsort(keys, kind="stable")
    assert (ko == keys[order]).all() and (vo == vals[order]).all()


def test_sort_ignores_bits_above_end_bit_and_handles_skew():
    rng = np.random.default_rng(9)
    n = 500_000
    keys = rng.integers(0, 1 << 40, n).astype(np.uint64)
    keys[: n // 2] = keys[0]                      # half the input in one bin
    keys |= np.uint64(1) << np.uint64(47)         # set in all keys: identical order whether ignored or not
    hi = (rng.integers(0, 2, n).astype(np.uint64)) << np.uint64(60)
    vals = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    ko, vo = _sort(keys | hi, vals, 45)
    order = np.argsort(keys & np.uint64((1 << 45) - 1), kind="stable")
    assert (vo == vals[order]).all() and (ko == (keys | hi)[order]).all()


def test_sort_all_equal_and_presorted_and_reversed():
    n = 70_000
    vals = np.arange(n, dtype=np.uint32)
    for keys in (np.full(n, 5 << 32, np.uint64), np.arange(n, dtype=np.uint64) << np.uint64(13),
                 (np.arange(n, dtype=np.uint64)[::-1].copy()) << np.uint64(13)):
        ko, vo = _sort(keys, vals, 45)
        order = np.argsort(keys, kind="stable")
        assert (ko == keys[order]).all() and (vo == vals[order]).all()


@pytest.mark.parametrize("n", [1, 7, 2047, 2048, 2049, 8191, 8192, 8193, 65536, 1_110_210, 5_000_003])
def test_inclusive_scan(n):
    from hugs_b200 import _native as N

    rng = np.random.default_rng(n)
    x = rng.integers(0, 40, n).astype(np.uint32)
    pad = (-n) % 4
    xin = torch.from_numpy(np.concatenate([x, np.zeros(pad, np.uint32)]).view(np.int32)).cuda()
    out = torch.empty_like(xin)
    tb = int(N.scan_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    rc = N.inclusive_scan(xin.data_ptr(), out.data_ptr(), n, temp.data_ptr(), tb,
                          C.c_void_p(torch.cuda.current_stream().cuda_stream))
    N.check(rc, "inclusive_scan")
    got = out.cpu().numpy().view(np.uint32)[:n]
    assert (got == np.cumsum(x, dtype=np.uint64).astype(np.uint32)).all()


def test_scan_rejects_small_temp():
    from hugs_b200 import _native as N

    x = torch.zeros(4096, dtype=torch.int32, device="cuda")
    temp = torch.empty(8, dtype=torch.uint8, device="cuda")
    rc = N.inclusive_scan(x.data_ptr(), x.data_ptr(), 4096, temp.data_ptr(), 8, None)
    assert rc != 0 and b"temp storage too small" in N.last_error()


@pytest.mark.parametrize("n,begin_bit,end_bit", [(100_003, 32, 45), (1_000_001, 32, 41), (4097, 32, 33), (50_000, 8, 27)])
def test_sort_pairs_range_is_a_stable_sort_on_those_bits_only(n, begin_bit, end_bit):
    """Second level of the rasterizer's binning: a stable sort on the tile bits alone of pairs that
    arrive in (depth, index) order reproduces the full (tile | depth) key order."""
    from hugs_b200 import _native as N

    rng = np.random.default_rng(n + end_bit)
    keys = rng.integers(0, 1 << 62, n).astype(np.uint64)
    vals = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    k_in = torch.from_numpy(keys.view(np.int64)).cuda()
    v_in = torch.from_numpy(vals.view(np.int32)).cuda()
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    tb = int(N.sort_temp_bytes(n))
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    rc = N.sort_pairs_range(N.ptr(k_in), N.ptr(v_in), N.ptr(k_out), N.ptr(v_out), n, begin_bit, end_bit,
                            temp.data_ptr(), tb, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    N.check(rc, "sort_pairs_range")
    torch.cuda.synchronize()
    digit = (keys >> np.uint64(begin_bit)) & np.uint64((1 << (end_bit - begin_bit)) - 1)
    order = np.argsort(digit, kind="stable")
    assert (k_out.cpu().numpy().view(np.uint64) == keys[order]).all()
    assert (v_out.cpu().numpy().view(np.uint32) == vals[order]).all()
