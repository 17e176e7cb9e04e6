"""GPU parity tests of the C-ABI pipeline (libb2d.so) against the checkers:
oracle/liboracle.so (restated reference arithmetic), oracle/libb2dtwin.so
(stage-for-stage CPU twin of the kernels) and the committed golden vectors.
All calls go through the C ABI (ctypes)."""
import ctypes
import zlib

import numpy as np
import pytest

import synthdata
from helpers import golden, oracle, orc_inflate, twin_deflate

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from zlib_ng_b200 import _native
    return _native


@pytest.fixture(scope="module")
def ctx(nat):
    c = nat.Context(0, 8 << 20)
    yield c
    c.close()


def deflate_host(nat, ctx, data: bytes, level: int, zdict: bytes = b"", final: bool = False, want_crc=True):
    L = nat.lib()
    cap = int(L.b2d_deflate_bound(len(data))) + 16
    out = ctypes.create_string_buffer(cap)
    n = ctypes.c_uint64(0)
    crc = ctypes.c_uint32(0)
    adler = ctypes.c_uint32(1)
    rc = L.b2d_deflate_host(ctx.handle, level, 15, data, len(data), zdict, len(zdict), 1 if final else 0,
                            ctypes.cast(out, ctypes.c_void_p), cap, ctypes.byref(n), ctypes.byref(crc),
                            ctypes.byref(adler))
    nat.check(rc)
    return out.raw[:n.value], crc.value, adler.value


CORPORA = [("fastq", 1_000_000, 5), ("mixed", 3_000_000, 6), ("fastq", 131072, 7), ("mixed", 131073, 8),
           ("fastq", 32768, 9), ("fastq", 32769, 10), ("mixed", 1023, 11), ("mixed", 1, 12)]


@pytest.mark.parametrize("kind,n,seed", CORPORA)
@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_matches_twin_and_round_trips(nat, ctx, kind, n, seed, level):
    data = synthdata.make(kind, n, seed).tobytes()
    out, crc, adler = deflate_host(nat, ctx, data, level)
    assert out[-4:] == b"\x00\x00\xff\xff"
    # reference-equivalent decoders: our restated inflate and stdlib zlib
    assert orc_inflate(out, len(data), need_final=False) == data
    assert zlib.decompressobj(-15).decompress(out) == data
    L = oracle()
    assert crc == L.orc_crc32(0, data, len(data))
    assert adler == L.orc_adler32(1, data, len(data))
    # stage-for-stage twin: identical bytes
    assert out == twin_deflate(data, level)


def test_dictionary_priming(nat, ctx):
    data = synthdata.make("fastq", 600_000, 21).tobytes()
    for cut in (100, 32768, 50_000, 200_000):
        zdict, body = data[:cut], data[cut:]
        out, crc, _ = deflate_host(nat, ctx, body, 6, zdict)
        assert zlib.decompressobj(-15, zdict=zdict[-32768:]).decompress(out) == body
        assert out == twin_deflate(body, 6, zdict)
        assert crc == zlib.crc32(body)
    # the dictionary must help (~0.8 % on FASTQ)
    a, _, _ = deflate_host(nat, ctx, data[200_000:331_072], 6, data[:200_000])
    b, _, _ = deflate_host(nat, ctx, data[200_000:331_072], 6)
    assert len(a) < len(b)


def test_batch_size_does_not_change_output(nat):
    data = synthdata.make("mixed", 5_000_000, 33).tobytes()
    outs = []
    for mb in (1 << 20, 4 << 20, 16 << 20):
        c = nat.Context(0, mb)
        outs.append(deflate_host(nat, c, data, 6)[0])
        c.close()
    assert outs[0] == outs[1] == outs[2]
    assert outs[0] == twin_deflate(data, 6)


@pytest.mark.parametrize("level", [1, 6])
def test_blocks_are_independent_given_their_dictionary(nat, ctx, level):
    """A 128 KiB block's bytes depend only on the block and the 32 KiB before it: compressing the blocks one
    call at a time, each with its predecessor's tail as zdict (what gzip_ng_threaded's workers do,
    gzip_ng_threaded.py:340-368), gives the same bytes as one call over the whole buffer."""
    data = synthdata.make("mixed", 1_000_000, 77).tobytes()
    whole = deflate_host(nat, ctx, data, level)[0]
    parts = []
    for off in range(0, len(data), 131072):
        parts.append(deflate_host(nat, ctx, data[off:off + 131072], level, zdict=data[max(0, off - 32768):off])[0])
    assert b"".join(parts) == whole
    assert twin_deflate(data, level) == whole


@pytest.mark.parametrize("strategy", [1, 2, 3, 4])
@pytest.mark.parametrize("level", [1, 6])
def test_strategies_match_twin(nat, ctx, level, strategy):
    """level | strategy << 8 (B2D_LEVEL_STRATEGY): Z_FILTERED, Z_HUFFMAN_ONLY, Z_RLE, Z_FIXED"""
    data = synthdata.make("mixed", 700_001, 21).tobytes() + b"\x07" * 70_000 + b"abcdefgh" * 9_000
    lv = level | (strategy << 8)
    out = deflate_host(nat, ctx, data, lv)[0]
    assert zlib.decompressobj(-15).decompress(out) == data
    assert out == twin_deflate(data, lv)
    if strategy == 4:
        assert (out[0] >> 1) & 3 in (0, 1)


def test_empty_and_final(nat, ctx):
    out, crc, adler = deflate_host(nat, ctx, b"", 6)
    assert out == b"\x00\x00\x00\xff\xff" and crc == 0 and adler == 1
    out, _, _ = deflate_host(nat, ctx, b"hello world" * 10, 6, final=True)
    assert out.endswith(b"\x03\x00")
    assert zlib.decompress(out, -15) == b"hello world" * 10


def test_incompressible_falls_back_to_stored(nat, ctx):
    rng = np.random.default_rng(5)
    data = rng.integers(0, 256, 400_000, dtype=np.uint8).tobytes()
    for level in (1, 3, 6):
        out, _, _ = deflate_host(nat, ctx, data, level)
        assert len(out) <= len(data) + 5 * 13 + 5 * 4 + 16
        assert zlib.decompressobj(-15).decompress(out) == data


def test_degenerate_inputs(nat, ctx):
    cases = [b"\0" * 300_000, b"a" * 70_001, b"ab" * 50_000, bytes(range(256)) * 1000,
             b"abcde" * 3 + b"x", b"aaaa", b"abc"]
    for data in cases:
        for level in (1, 6):
            out, crc, _ = deflate_host(nat, ctx, data, level)
            assert zlib.decompressobj(-15).decompress(out) == data
            assert out == twin_deflate(data, level)
            assert crc == zlib.crc32(data)


def test_output_buffer_too_small(nat, ctx):
    L = nat.lib()
    data = synthdata.make("fastq", 200_000, 3).tobytes()
    out = ctypes.create_string_buffer(1000)
    n = ctypes.c_uint64(0)
    crc = ctypes.c_uint32(0)
    rc = L.b2d_compress_and_crc(ctx.handle, 6, data, len(data), None, 0, ctypes.cast(out, ctypes.c_void_p), 1000,
                                ctypes.byref(n), ctypes.byref(crc))
    assert rc == nat.B2D_BUF_ERROR
    assert "exceeds buffer size" in nat.last_error()
    rc = L.b2d_compress_and_crc(ctx.handle, 12, data, len(data), None, 0, ctypes.cast(out, ctypes.c_void_p), 1000,
                                ctypes.byref(n), ctypes.byref(crc))
    assert rc == nat.B2D_STREAM_ERROR and "Bad compression level" in nat.last_error()


# ------------------------------------------------------------------ checksums
def _crc(nat, ctx, data, start=0):
    v = ctypes.c_uint32(0)
    nat.check(nat.lib().b2d_crc32(ctx.handle, start, data, len(data), ctypes.byref(v)))
    return v.value


def _adler(nat, ctx, data, start=1):
    v = ctypes.c_uint32(0)
    nat.check(nat.lib().b2d_adler32(ctx.handle, start, data, len(data), ctypes.byref(v)))
    return v.value


def test_checksums_and_deflate_over_many_pieces_pageable_and_pinned(nat):
    """Host buffers larger than one batch take the overlapped path (pinned staging slots for pageable
    memory, direct DMA for pinned): same values / bytes as one piece, for crc32, adler32 and deflate with
    either checksum."""
    data = synthdata.make("mixed", 5_300_001, 41).tobytes()
    small = nat.Context(0, 1 << 20)
    big = nat.Context(0, 16 << 20)
    try:
        for start in (0, 0xDEADBEEF):
            assert _crc(nat, small, data, start) == zlib.crc32(data, start)
            assert _adler(nat, small, data, start) == zlib.adler32(data, start)
        want = deflate_host(nat, big, data, 6, final=True)
        got = deflate_host(nat, small, data, 6, final=True)
        assert got == want
        assert got[1] == zlib.crc32(data) and got[2] == zlib.adler32(data)
        # pinned source
        p = small.host_alloc(len(data))
        ctypes.memmove(p, data, len(data))
        v = ctypes.c_uint32(0)
        nat.check(nat.lib().b2d_crc32(small.handle, 7, ctypes.c_void_p(p), len(data), ctypes.byref(v)))
        assert v.value == zlib.crc32(data, 7)
        nat.check(nat.lib().b2d_adler32(small.handle, 7, ctypes.c_void_p(p), len(data), ctypes.byref(v)))
        assert v.value == zlib.adler32(data, 7)
        small.host_free(p)
    finally:
        small.close()
        big.close()


def test_checksum_kat_rows(nat, ctx):
    from helpers import b64
    kat = golden("checksum_kat.json")
    for r in kat["crc32"]:
        if r["null"]:
            continue
        data = b64(r["data"])
        assert _crc(nat, ctx, data, r["start"]) == r["expect"], r
    for r in kat["adler32"]:
        if r["null"]:
            continue
        data = b64(r["data"])
        assert _adler(nat, ctx, data, r["start"]) == r["expect"], r
    for s, start, expect in kat["python"]["crc32"]:
        assert _crc(nat, ctx, s.encode(), start) == expect
    for s, start, expect in kat["python"]["adler32"]:
        assert _adler(nat, ctx, s.encode(), start) == expect


def test_checksums_vs_reference_vectors(nat, ctx):
    for r in golden("reference_vectors.json")["checksums"]:
        data = synthdata.make(r["kind"], r["n"], r["seed"]).tobytes()
        assert _crc(nat, ctx, data, r["start"]) == r["crc32"]
        assert _adler(nat, ctx, data, r["start"]) == r["adler32"]
    L = nat.lib()
    for r in golden("reference_vectors.json")["combine"]:
        assert L.b2d_crc32_combine(r["crc1"], r["crc2"], r["len2"]) == r["expect"]


def test_checksum_sizes_and_starts(nat, ctx):
    # tests/test_compat.py:26-32,66-77 in the reference: sizes x start values vs an independent implementation
    rng = np.random.default_rng(17)
    L = oracle()
    sizes = [0, 1, 8, 15, 16, 17, 63, 64, 4095, 5552, 5553, 16383, 16384, 16385, 16400, 65536, 131071,
             524288, 1 << 20, (1 << 20) + 12345, 9_000_001]
    starts = [0, 1, 0xFFFFFFFF, 0xDEADBEEF, 65520, 65521, 0xFFF1FFF1, 123456789]
    for n in sizes:
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        for s in starts:
            assert _crc(nat, ctx, data, s) == L.orc_crc32(s, data, n), (n, s)
            assert _adler(nat, ctx, data, s) == L.orc_adler32(s, data, n), (n, s)
        assert _crc(nat, ctx, data) == zlib.crc32(data)
        assert _adler(nat, ctx, data) == zlib.adler32(data)


def test_checksum_split_and_combine(nat, ctx):
    data = synthdata.make("mixed", 6_000_000, 2).tobytes()
    L = nat.lib()
    cuts = [0, 1, 700_001, 1_500_000, 1_500_016, 3_000_000, 4_444_444, 5_999_999, 6_000_000]
    crc, adler = 0, 1
    for a, b in zip(cuts, cuts[1:]):
        part = data[a:b]
        crc = L.b2d_crc32_combine(crc, _crc(nat, ctx, part), len(part))
        adler = L.b2d_adler32_combine(adler, _adler(nat, ctx, part), len(part))
    assert crc == zlib.crc32(data) and adler == zlib.adler32(data)


def test_device_resident_variants(nat, ctx):
    L = nat.lib()
    data = synthdata.make("fastq", 3_000_000, 41).tobytes()
    zdict = synthdata.make("fastq", 32768, 42).tobytes()
    n = len(data)
    d_in = ctx.device_alloc(len(zdict) + n + 64)
    cap = int(L.b2d_deflate_bound(n))
    cap = (cap + 3) & ~3
    d_out = ctx.device_alloc(cap)
    try:
        ctx.h2d(d_in, zdict + data, len(zdict) + n)
        out_n = ctypes.c_uint64(0)
        crc = ctypes.c_uint32(0)
        ms = ctypes.c_float(0)
        nat.check(L.b2d_deflate_device(ctx.handle, 6, ctypes.c_void_p(d_in), len(zdict), n, ctypes.c_void_p(d_out), cap,
                                       ctypes.byref(out_n), ctypes.byref(crc), ctypes.byref(ms)))
        host = ctypes.create_string_buffer(out_n.value)
        ctx.d2h(host, d_out, out_n.value)
        assert host.raw == twin_deflate(data, 6, zdict)
        assert zlib.decompressobj(-15, zdict=zdict).decompress(host.raw) == data
        assert crc.value == zlib.crc32(data) and ms.value > 0
        v = ctypes.c_uint32(0)
        nat.check(L.b2d_crc32_device(ctx.handle, 7, ctypes.c_void_p(d_in + 5), n, ctypes.byref(v), None))
        assert v.value == zlib.crc32((zdict + data)[5:5 + n], 7)
        nat.check(L.b2d_adler32_device(ctx.handle, 7, ctypes.c_void_p(d_in + 3), n, ctypes.byref(v), None))
        assert v.value == zlib.adler32((zdict + data)[3:3 + n], 7)
    finally:
        ctx.device_free(d_in)
        ctx.device_free(d_out)
    assert ctx.launch_count() > 0


def test_pinned_multi_batch_pipelined_path(nat):
    """pinned in/out and more than one batch -> the overlapped H2D / kernels / D2H path; same bytes as ever"""
    L = nat.lib()
    data = synthdata.make("mixed", 9_000_001, 55).tobytes()
    zdict = synthdata.make("fastq", 20_000, 56).tobytes()
    n = len(data)
    c = nat.Context(0, 1 << 20)
    cap = int(L.b2d_deflate_bound(n)) + 16
    h_in, h_out = c.host_alloc(n), c.host_alloc(cap)
    try:
        ctypes.memmove(h_in, data, n)
        for final in (0, 1):
            out_n, crc = ctypes.c_uint64(0), ctypes.c_uint32(0)
            nat.check(L.b2d_deflate_host(c.handle, 6, 15, ctypes.c_void_p(h_in), n, zdict, len(zdict), final,
                                         ctypes.c_void_p(h_out), cap, ctypes.byref(out_n), ctypes.byref(crc), None))
            got = ctypes.string_at(h_out, out_n.value)
            want = twin_deflate(data, 6, zdict) + (b"\x03\x00" if final else b"")
            assert got == want
            assert crc.value == zlib.crc32(data)
        # too small an output buffer is reported, not overrun
        out_n, crc = ctypes.c_uint64(0), ctypes.c_uint32(0)
        rc = L.b2d_deflate_host(c.handle, 6, 15, ctypes.c_void_p(h_in), n, None, 0, 0, ctypes.c_void_p(h_out), 2_000_000,
                                ctypes.byref(out_n), ctypes.byref(crc), None)
        assert rc == nat.B2D_BUF_ERROR
    finally:
        c.host_free(h_in)
        c.host_free(h_out)
        c.close()
