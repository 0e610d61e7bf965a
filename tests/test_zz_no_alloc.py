"""fqlpop_philox_fill_noalloc (include/fqlpop.h): the allocation-free variant of fqlpop_philox_fill, whose device staging
buffer is a fixed file-scope device array. Called through the C-ABI and checked against the allocating entry point on two
calls in a row (the second one smaller, with other arguments: the shared array keeps the first call's contents and none of
them may leak into the result). Collected last on purpose (zz)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from fqlpop._lib import lib
    u64, i32, u32, vp = ctypes.c_uint64, ctypes.c_int32, ctypes.c_uint32, ctypes.c_void_p
    for name in ("fqlpop_philox_fill", "fqlpop_philox_fill_noalloc"):
        f = getattr(lib, name)
        f.argtypes = [u64, u64, i32, i32, i32, u32, vp]
        f.restype = ctypes.c_int
    lib.fqlpop_philox_fill_noalloc_max_words.argtypes = []
    lib.fqlpop_philox_fill_noalloc_max_words.restype = ctypes.c_int32
    lib.fqlpop_last_error.restype = ctypes.c_char_p
    return lib


def _fill(lib, name, seed, step, rows, slots, kind, n=0):
    shape = (rows, slots, 4) if kind == 0 else (rows, slots)
    out = np.full(shape, 0xA5A5A5A5 if kind in (0, 2) else np.nan, np.uint32 if kind in (0, 2) else np.float32)
    rc = getattr(lib, name)(seed, step, rows, slots, kind, n, out.ctypes.data)
    assert rc == 0, lib.fqlpop_last_error().decode()
    return out


def test_no_alloc_philox_fill_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    lib = _lib()
    max_words = lib.fqlpop_philox_fill_noalloc_max_words()
    assert max_words == 262144
    # first call: all four kinds at the full capacity of the array (kind 0 writes 4 words per draw)
    calls = [(0x1234567890ABCDEF, 77, 256, 256, 0, 0), (0x1234567890ABCDEF, 77, 512, 512, 1, 0),
             (3, 2**40 + 5, 512, 512, 2, 1_000_003), (2**63 + 11, 9, 512, 512, 3, 0),
             # second round: smaller, other seeds / shapes -- stale words of the first must not show up
             (99, 1, 257, 23, 0, 0), (99, 1, 257, 23, 1, 0), (7, 123456789, 31, 5, 2, 17), (5, 2, 129, 3, 3, 0)]
    for seed, step, rows, slots, kind, n in calls:
        got = _fill(lib, "fqlpop_philox_fill_noalloc", seed, step, rows, slots, kind, n)
        want = _fill(lib, "fqlpop_philox_fill", seed, step, rows, slots, kind, n)
        if kind in (0, 2):
            assert np.array_equal(got, want)          # integer work: bit-exact
        else:
            assert np.isfinite(got).all()
            # same tolerance as the oracle check of the stream (tests/test_rollout_gpu.py); it is the same kernel
            assert np.allclose(got, want, atol=2e-6, rtol=1e-5)
    # the very same call twice in a row gives the very same words
    a = _fill(lib, "fqlpop_philox_fill_noalloc", 42, 7, 300, 40, 3)
    b = _fill(lib, "fqlpop_philox_fill_noalloc", 42, 7, 300, 40, 3)
    assert np.array_equal(a, b)


def test_no_alloc_philox_fill_rejects_requests_beyond_the_fixed_scratch():
    lib = _lib()
    max_words = lib.fqlpop_philox_fill_noalloc_max_words()
    out = np.zeros(8, np.uint32)   # never written: the call must fail before touching the GPU
    assert lib.fqlpop_philox_fill_noalloc(1, 1, max_words + 1, 1, 1, 0, out.ctypes.data) != 0
    assert "exceed" in lib.fqlpop_last_error().decode()
    assert lib.fqlpop_philox_fill_noalloc(1, 1, max_words // 4 + 1, 1, 0, 0, out.ctypes.data) != 0   # kind 0: 4 words per draw
    assert lib.fqlpop_philox_fill_noalloc(1, 1, 2**20, 2**20, 1, 0, out.ctypes.data) != 0            # no 32-bit overflow
    assert not out.any()
