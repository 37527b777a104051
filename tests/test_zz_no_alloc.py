"""sb200_prune_rows_static: the allocation-free variant of sb200_prune_rows (include/surfel_b200.h), whose block counts
live in a file-scope `static __device__` array of the library instead of caller scratch.  It has no counterpart in the
reference; the checker here is the existing entry point, and the bar is
the one tests/test_churn.py holds the compaction to: byte-exact.

Collected last (tests/conftest.py) so that under `pytest -x` a failure here cannot hide another test.  Everything goes
through the C ABI with ctypes; torch only provides the device buffers."""
import ctypes

import pytest
import torch

import util

pytestmark = pytest.mark.gpu

MAX_ROWS = 16 * 1024 * 1024      # SB200_PRUNE_STATIC_MAX_ROWS
# bytes per row: xyz, rgb, opacity, scaling, rotation, an int64 id, and two widths that take the byte-copy path
ROW_BYTES = (12, 12, 4, 8, 16, 8, 1, 3)


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _arrays(P, seed, device, row_bytes=ROW_BYTES):
    g = torch.Generator(device="cpu").manual_seed(seed)
    src = [torch.randint(0, 256, (max(P, 1), nb), dtype=torch.uint8, generator=g).to(device)[:P] for nb in row_bytes]
    mask = (torch.rand(max(P, 1), generator=g) < (seed % 7) / 7.0)[:P].to(device)
    return src, mask


def _c_args(src, dst):
    n = len(src)
    sp = (ctypes.c_void_p * n)(*[s.data_ptr() for s in src])
    dp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in dst])
    rb = (ctypes.c_int * n)(*[t.shape[1] for t in src])
    return n, sp, dp, rb


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _existing(lib, P, src, mask, device):
    """The existing entry point: caller scratch, host read-back of the count."""
    dst = [torch.full((max(P, 1), nb), 0xAB, dtype=torch.uint8, device=device) for nb in (t.shape[1] for t in src)]
    scratch = torch.empty(int(lib.sb200_prune_scratch_bytes(P)) + 256, dtype=torch.uint8, device=device)
    kept = ctypes.c_int(-1)
    n, sp, dp, rb = _c_args(src, dst)
    rc = lib.sb200_prune_rows(P, _vp(mask), n, sp, dp, rb, _vp(scratch), ctypes.byref(kept), _stream(device))
    assert rc == 0, lib.sb200_last_error()
    torch.cuda.synchronize(device)
    return dst, kept.value


def _static(lib, P, src, mask, device, check_no_alloc=False):
    """The new entry point: no scratch argument, the count stays on the device."""
    dst = [torch.full((max(P, 1), nb), 0xAB, dtype=torch.uint8, device=device) for nb in (t.shape[1] for t in src)]
    kept = torch.full((1,), -1, dtype=torch.int32, device=device)
    n, sp, dp, rb = _c_args(src, dst)
    stream = _stream(device)
    torch.cuda.synchronize(device)
    free0 = torch.cuda.mem_get_info(device)[0]
    rc = lib.sb200_prune_rows_static(P, _vp(mask), n, sp, dp, rb, _vp(kept), stream)
    assert rc == 0, lib.sb200_last_error()
    torch.cuda.synchronize(device)
    free1 = torch.cuda.mem_get_info(device)[0]
    if check_no_alloc:   # driver-level free memory: the call left no allocation behind (first calls load kernels)
        assert free1 >= free0, (free0, free1)
    return dst, int(kept.item())


def _compare(lib, P, seed, device, check_no_alloc=False, row_bytes=ROW_BYTES):
    src, mask = _arrays(P, seed, device, row_bytes)
    want, want_kept = _existing(lib, P, src, mask, device)
    got, got_kept = _static(lib, P, src, mask, device, check_no_alloc)
    assert got_kept == want_kept == int((~mask).sum())
    for a, (g, w) in enumerate(zip(got, want)):
        assert torch.equal(g, w), ("array", a, "P", P)                 # kept rows AND the untouched 0xAB tail
        assert torch.equal(g[:got_kept], src[a][~mask]), ("array", a)  # and both equal torch's x[~mask]


def test_static_scratch_prune_equals_prune_rows_on_two_calls_in_a_row():
    lib = util.load_mine()._load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        # a large call, then smaller ones that find the previous call's counts still in the array
        _compare(lib, 300_000, 3, dev)
        _compare(lib, 5_001, 5, dev, check_no_alloc=True)
        _compare(lib, 1_024, 6, dev, check_no_alloc=True)
        _compare(lib, 1, 1, dev, check_no_alloc=True)
        _compare(lib, 2_000_000, 2, dev, check_no_alloc=True)
        _compare(lib, 0, 4, dev)


def test_static_scratch_prune_at_its_maximum_size():
    lib = util.load_mine()._load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        _compare(lib, MAX_ROWS, 3, dev, row_bytes=(4, 1))   # every word of the static array in use
        _compare(lib, 70_000, 2, dev, check_no_alloc=True)


def test_static_scratch_prune_rejects_sizes_beyond_its_maximum():
    lib = util.load_mine()._load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        src, mask = _arrays(8, 1, dev)
        dst = [torch.zeros_like(s) for s in src]
        kept = torch.full((1,), -1, dtype=torch.int32, device=dev)
        n, sp, dp, rb = _c_args(src, dst)
        rc = lib.sb200_prune_rows_static(MAX_ROWS + 1, _vp(mask), n, sp, dp, rb, _vp(kept), _stream(dev))
        assert rc != 0
        msg = lib.sb200_last_error().decode()
        assert "SB200_PRUNE_STATIC_MAX_ROWS" in msg and str(MAX_ROWS) in msg, msg
        torch.cuda.synchronize(dev)
        assert int(kept.item()) == -1 and all(int(d.sum()) == 0 for d in dst)   # nothing was launched
        assert lib.sb200_prune_rows_static(8, _vp(mask), n, sp, dp, rb, None, _stream(dev)) != 0


def test_static_scratch_address_is_taken_per_device():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lib = util.load_mine()._load()
    for d in (1, 0, 1):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            _compare(lib, 40_000 + d, 3 + d, dev)
