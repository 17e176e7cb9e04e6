"""GPU test of b2d_adler32_device_noalloc (include/b2d.h): the allocation-free Adler-32 entry point, whose scratch is a
file-scope device array instead of a cudaMalloc'ed workspace.  It has no counterpart in the reference,
so the checker is the existing entry point b2d_adler32_device — bit-exact, as for
every checksum test here — and stdlib zlib.adler32 beside it.

Collected last on purpose (test_zz_*): a failure here must not hide other tests.
Passes alone in a fresh process and after the other tests (the array keeps the last call's contents: two calls in a row,
the second over less data, cover that)."""
import ctypes
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from zlib_ng_b200 import _native
    return _native


def test_noalloc_adler32_matches_the_existing_entry_point_twice_in_a_row(nat):
    L = nat.lib()
    ctx = nat.Context(0, 1 << 20)            # only for the test's own device buffer and the existing entry point
    n_max = 48 * 1024 * 1024 + 12345
    rng = np.random.default_rng(20261105)
    data = rng.integers(0, 256, n_max, dtype=np.uint8)
    d = ctx.device_alloc(n_max + 64)
    try:
        ctx.h2d(d, data.ctypes.data_as(ctypes.c_void_p), n_max)
        # two calls in a row on the shared scratch: a big one (many partial sums), then a smaller, misaligned one
        # that must not see what the first left behind; then sizes around the kernels' edges
        cases = [(0, n_max, 1), (3, 5 * 1024 * 1024 + 7, 0xDEADBEEF), (0, 1, 1), (5, 15, 7), (16, 16, 1), (1, 16384 * 3 + 31, 65520),
                 (0, 0, 432), (0, n_max, 1)]
        for off, n, start in cases:
            want = ctypes.c_uint32(0)
            nat.check(L.b2d_adler32_device(ctx.handle, start, ctypes.c_void_p(d + off), n, ctypes.byref(want), None))
            got = ctypes.c_uint32(0xFFFFFFFF)
            nat.check(L.b2d_adler32_device_noalloc(0, start, ctypes.c_void_p(d + off), n, ctypes.byref(got)))
            assert got.value == want.value, (off, n, start)
            assert got.value == zlib.adler32(data[off:off + n].tobytes(), start), (off, n, start)
    finally:
        ctx.device_free(d)
        ctx.close()


def test_noalloc_rejects_sizes_beyond_its_maximum(nat):
    L = nat.lib()
    limit = int(L.b2d_noalloc_max_bytes())
    assert limit == 1 << 30
    out = ctypes.c_uint32(0)
    # refused before anything is launched: the pointer is never read
    rc = L.b2d_adler32_device_noalloc(0, 1, ctypes.c_void_p(256), limit + 1, ctypes.byref(out))
    assert rc == nat.B2D_STREAM_ERROR
    msg = nat.last_error()
    assert "exceed the maximum" in msg and str(limit) in msg
    assert L.b2d_adler32_device_noalloc(0, 1, None, 16, ctypes.byref(out)) == nat.B2D_STREAM_ERROR
    assert L.b2d_adler32_device_noalloc(9999, 1, ctypes.c_void_p(256), 16, ctypes.byref(out)) != 0
