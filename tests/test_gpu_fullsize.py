"""GPU parity at BASELINE.json's full sizes, through size-independent properties:

* checksums of a 4 GiB device-resident buffer equal the checker's value of the same bytes, for two start values,
  and equal the fold of eight unequal shards with crc32_combine / adler32_combine ("a checksum of checksums");
* a buffer longer than 2^32 bytes goes through the public zlib_ng.crc32 / adler32 (the reference loops in
  UINT32_MAX slices there: zlib_ngmodule.c:1469-1500, 1531-1562; its test: tests/test_zlib_compliance.py:131-137);
* level 1 over 1 GiB FASTQ-like and level 6 over 4 GiB mixed-entropy data, device-resident, 128 KiB blocks:
  the whole stream inflates back to the input (stdlib zlib here, the decoder family the reference's own tests
  cross-check with), the CRC returned by the device equals the checker's, and the size is sane.

All calls go through the C ABI.  The checkers (stdlib zlib, oracle/liboracle.so) run on the host in seconds.
"""
import ctypes
import zlib

import numpy as np
import pytest

import synthdata
from helpers import oracle

pytestmark = pytest.mark.gpu

GIB = 1 << 30
MIB = 1 << 20


@pytest.fixture(scope="module")
def nat():
    from zlib_ng_b200 import _native
    return _native


@pytest.fixture(scope="module")
def ctx(nat):
    c = nat.Context(0, 256 * MIB)
    yield c
    c.close()


def _tiled(kind, n, seed, base_mib=256):
    base = synthdata.make(kind, min(n, base_mib * MIB), seed)
    return np.ascontiguousarray(np.tile(base, (n + base.size - 1) // base.size)[:n])


def _host_crc(data: np.ndarray, start=0):
    """zlib.crc32 of a (possibly > 4 GiB) array, in 1 GiB pieces."""
    v = start
    for off in range(0, data.size, GIB):
        v = zlib.crc32(memoryview(data[off:off + GIB]), v)
    return v


def _host_adler(data: np.ndarray, start=1):
    v = start
    for off in range(0, data.size, GIB):
        v = zlib.adler32(memoryview(data[off:off + GIB]), v)
    return v


def _dev_sum(nat, ctx, fn, start, d_ptr, n):
    v, ms = ctypes.c_uint32(0), ctypes.c_float(0)
    nat.check(fn(ctx.handle, start, ctypes.c_void_p(d_ptr), n, ctypes.byref(v), ctypes.byref(ms)))
    return v.value


def test_checksums_4gib_device_resident(nat, ctx):
    L = nat.lib()
    n = 4 * GIB
    data = _tiled("mixed", n, 2)
    d_in = ctx.device_alloc(n + 64)
    try:
        ctx.h2d(d_in, data.ctypes.data_as(ctypes.c_void_p), n)
        O = oracle()
        for start in (0, 0xDEADBEEF):
            want = _host_crc(data, start)
            assert _dev_sum(nat, ctx, L.b2d_crc32_device, start, d_in, n) == want
        # the restated reference arithmetic agrees with stdlib on a 64 MiB slice of the same buffer
        head = data[:64 * MIB].tobytes()
        assert O.orc_crc32(0, head, len(head)) == zlib.crc32(head)
        for start in (1, 0xDEADBEEF):
            assert _dev_sum(nat, ctx, L.b2d_adler32_device, start, d_in, n) == _host_adler(data, start)
        # eight unequal shards (what eight GPUs would hold), folded on the host: SURVEY §8(e)
        cuts = [0, 300 * MIB + 1, 700 * MIB + 13, GIB + 5, 2 * GIB - 7, 2 * GIB + 1234567, 3 * GIB + 1, 4 * GIB - 99, n]
        crc, adler = 0, 1
        for a, b in zip(cuts, cuts[1:]):
            crc = L.b2d_crc32_combine(crc, _dev_sum(nat, ctx, L.b2d_crc32_device, 0, d_in + a, b - a), b - a)
            adler = L.b2d_adler32_combine(adler, _dev_sum(nat, ctx, L.b2d_adler32_device, 1, d_in + a, b - a), b - a)
        assert crc == _host_crc(data) and adler == _host_adler(data)
    finally:
        ctx.device_free(d_in)


def test_checksums_of_more_than_4gib_through_the_python_api():
    from zlib_ng_b200 import zlib_ng
    n = 4 * GIB + 5 * MIB + 3
    data = _tiled("mixed", n, 4, base_mib=64)
    assert zlib_ng.crc32(memoryview(data)) == _host_crc(data)
    assert zlib_ng.adler32(memoryview(data)) == _host_adler(data)
    assert zlib_ng.crc32(memoryview(data), 0xDEADBEEF) == _host_crc(data, 0xDEADBEEF)


def _deflate_device_and_check(nat, ctx, data: np.ndarray, level: int, max_ratio: float):
    L = nat.lib()
    n = data.size
    d_in = ctx.device_alloc(n + 64)
    cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
    d_out = ctx.device_alloc(cap)
    try:
        ctx.h2d(d_in, data.ctypes.data_as(ctypes.c_void_p), n)
        out_n, crc, ms = ctypes.c_uint64(0), ctypes.c_uint32(0), ctypes.c_float(0)
        nat.check(L.b2d_deflate_device(ctx.handle, level, ctypes.c_void_p(d_in), 0, n, ctypes.c_void_p(d_out), cap,
                                       ctypes.byref(out_n), ctypes.byref(crc), ctypes.byref(ms)))
        assert crc.value == _host_crc(data)
        assert 0 < out_n.value <= max_ratio * n
        comp = np.empty(out_n.value, dtype=np.uint8)
        ctx.d2h(comp.ctypes.data_as(ctypes.c_void_p), d_out, out_n.value)
    finally:
        ctx.device_free(d_in)
        ctx.device_free(d_out)
    # inflate the whole raw stream piecewise and compare as it comes (no second 4 GiB copy)
    d = zlib.decompressobj(-15)
    pos = 0
    piece = 32 * MIB
    for off in range(0, comp.size, piece):
        back = d.decompress(memoryview(comp[off:off + piece]))
        assert back == memoryview(data[pos:pos + len(back)]), f"mismatch in output range starting at {pos}"
        pos += len(back)
    pos += len(d.flush())
    assert pos == n and d.unused_data == b""
    return out_n.value / n


def _reference_ratio(kind, seed, level):
    """the reference's own size on the 256 MiB segment these workloads tile (tests/golden/fullsize_ratios.json, written
    by tools/make_golden_fullsize.py from oracle/_ref: 1 MiB blocks through _ParallelCompress, as its threaded writer)"""
    import json
    import os
    rows = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_ratios.json")))["rows"]
    (row,) = [r for r in rows if (r["kind"], r["seed"], r["level"]) == (kind, seed, level)]
    assert row["base_bytes"] == 256 * MIB
    return row["reference_ratio"]


def test_level1_1gib_fastq_device_resident(nat, ctx):
    # BASELINE.json configs[2]; north star: compressed size within 3 % of the reference's at the same level (here a third smaller)
    ratio = _deflate_device_and_check(nat, ctx, _tiled("fastq", GIB, 1), 1, 0.60)
    assert ratio <= _reference_ratio("fastq", 1, 1) * 1.03


def test_level6_4gib_mixed_device_resident(nat, ctx):
    # BASELINE.json configs[3]; tolerance 1.03x of the reference's size
    ratio = _deflate_device_and_check(nat, ctx, _tiled("mixed", 4 * GIB, 2), 6, 0.60)
    assert ratio <= _reference_ratio("mixed", 2, 6) * 1.03
