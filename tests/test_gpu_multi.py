"""GPU tests of the multi-device entry points (b2d_mctx_*, b2d_compress_and_crc_multi, b2d_deflate_host_multi):
ONE stream cut over several GPU contexts, outputs concatenated in order, CRCs folded with crc32_combine — what the
reference's N worker threads + writer thread do on one stream (src/zlib_ng/gzip_ng_threaded.py:316-321, 382-398).

A box with one GPU still exercises all of it: a multi-device context may name device 0 several times (one context
per entry).  Checks: the bytes equal the single-context call's (blocks are independent given the raw input, so this
is an equality), the reference's own decoder (oracle/_ref: zlib_ng.decompressobj) gives the input
back, the CRC equals the reference's crc32, the block sizes add up."""
import ctypes
import zlib

import numpy as np
import pytest

import synthdata
from helpers import reference

pytestmark = pytest.mark.gpu
BLOCK = 131072
MIB = 1 << 20


@pytest.fixture(scope="module")
def nat():
    from zlib_ng_b200 import _native
    return _native


@pytest.fixture(scope="module")
def single(nat):
    c = nat.Context(0, 8 * MIB)
    yield c
    c.close()


def ref_inflate(stream: bytes, zdict: bytes = b"") -> bytes:
    """raw-DEFLATE bytes -> input, through the REFERENCE's decoder when oracle/_ref is there (it travels to the GPU box)"""
    ref = reference()
    mod = ref if ref is not None else zlib
    d = mod.decompressobj(-15, zdict=zdict) if zdict else mod.decompressobj(-15)
    return d.decompress(stream + b"\x03\x00") + d.flush()


def ref_crc32(data: bytes) -> int:
    ref = reference()
    return (ref if ref is not None else zlib).crc32(data)


def single_call(nat, ctx, data, level, zdict=b""):
    L = nat.lib()
    cap = int(L.b2d_deflate_bound(len(data))) + 64
    out = ctypes.create_string_buffer(cap)
    out_n, crc = ctypes.c_uint64(0), ctypes.c_uint32(0)
    nb = (len(data) + BLOCK - 1) // BLOCK
    offs, cnt = (ctypes.c_uint64 * (nb + 1))(), ctypes.c_uint64(0)
    nat.check(L.b2d_compress_and_crc_indexed(ctx.handle, level, data, len(data), zdict if zdict else None, len(zdict), out, cap,
                                             ctypes.byref(out_n), ctypes.byref(crc), offs, nb + 1, ctypes.byref(cnt)))
    sizes = [offs[i + 1] - offs[i] for i in range(nb)]
    return out.raw[:out_n.value], crc.value, sizes


def multi_call(nat, m, data_addr, n, level, zdict=b"", cap=None, out_addr=None):
    L = nat.lib()
    cap = cap if cap is not None else int(L.b2d_deflate_bound(n)) + 64
    out = None
    if out_addr is None:
        out = ctypes.create_string_buffer(max(cap, 1))
        out_addr = ctypes.addressof(out)
    got, crc, sizes = m.compress_and_crc(level, data_addr, n, zdict if zdict else None, len(zdict), out_addr, cap, want_sizes=True)
    return ctypes.string_at(out_addr, got), crc, sizes


@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 0, 0]])
@pytest.mark.parametrize("level", [1, 6])
def test_one_stream_over_several_contexts_equals_single_context(nat, single, devices, level):
    data = synthdata.make("mixed", 41 * MIB + 12345, 31).tobytes()        # ragged end, 6 batches of 8 MiB
    zdict = synthdata.make("fastq", 40_000, 32).tobytes()
    want, want_crc, want_sizes = single_call(nat, single, data, level, zdict)
    m = nat.MultiContext(devices, 8 * MIB)
    try:
        got, crc, sizes = multi_call(nat, m, data, len(data), level, zdict)
    finally:
        m.close()
    assert got == want                                  # byte-identical to the one-context stream
    assert crc == want_crc == ref_crc32(data)
    assert sizes == want_sizes and sum(sizes) == len(got)
    assert ref_inflate(got, zdict[-32768:]) == data     # the reference's decoder gives the input back


def test_pinned_buffers_and_every_block_starts_where_the_sizes_say(nat):
    L = nat.lib()
    n = 64 * MIB
    data = synthdata.make("fastq", n, 33)
    m = nat.MultiContext([0, 0], 8 * MIB)
    ctx = nat.Context(0, 1 * MIB)
    cap = int(L.b2d_deflate_bound(n)) + 64
    h_in, h_out = ctx.host_alloc(n), ctx.host_alloc(cap)
    try:
        ctypes.memmove(h_in, data.ctypes.data, n)
        got, crc, sizes = multi_call(nat, m, h_in, n, 6, cap=cap, out_addr=h_out)
        assert crc == zlib.crc32(data)
        assert ref_inflate(got) == data.tobytes()
        # the index is usable: each block, cut out by its size, inflates on its own given the 32 KiB before it
        raw = data.tobytes()
        off = 0
        for b in (0, 1, 63, 64, 65, len(sizes) - 1):            # around the first batch boundary too
            start = sum(sizes[:b])
            piece = got[start:start + sizes[b]]
            zd = raw[max(0, b * BLOCK - 32768):b * BLOCK]
            assert ref_inflate(piece, zd) == raw[b * BLOCK:(b + 1) * BLOCK]
        del off
    finally:
        ctx.host_free(h_in)
        ctx.host_free(h_out)
        ctx.close()
        m.close()


def test_small_inputs_empty_input_and_overflow(nat, single):
    m = nat.MultiContext([0, 0], 8 * MIB)
    try:
        for n in (0, 1, 1000, BLOCK, BLOCK + 1, 3 * MIB):
            data = synthdata.make("mixed", n, 40 + (n % 7)).tobytes() if n else b""
            want, want_crc, want_sizes = single_call(nat, single, data, 6)
            got, crc, sizes = multi_call(nat, m, data, n, 6)
            assert got == want and crc == want_crc and sizes == want_sizes
        data = np.random.default_rng(5).integers(0, 256, 20 * MIB, dtype=np.uint8).tobytes()
        with pytest.raises(nat.NativeError) as e:
            multi_call(nat, m, data, len(data), 6, cap=len(data) // 2)
        assert e.value.code == nat.B2D_BUF_ERROR
        # the contexts are still usable after a failed call
        got, crc, _ = multi_call(nat, m, data, len(data), 6)
        assert crc == zlib.crc32(data) and ref_inflate(got) == data
    finally:
        m.close()


def test_general_form_final_block_adler_and_small_window(nat):
    L = nat.lib()
    data = synthdata.make("mixed", 20 * MIB, 44).tobytes()
    m = nat.MultiContext([0, 0], 8 * MIB)
    try:
        cap = int(L.b2d_deflate_bound(len(data))) + 64
        out = ctypes.create_string_buffer(cap)
        out_n, crc, adler, cnt = ctypes.c_uint64(0), ctypes.c_uint32(0), ctypes.c_uint32(1), ctypes.c_uint64(0)
        nat.check(L.b2d_deflate_host_multi(m.handle, 6, 12, data, len(data), None, 0, 1, out, cap, ctypes.byref(out_n),
                                           ctypes.byref(crc), ctypes.byref(adler), None, 0, ctypes.byref(cnt)))
        stream = out.raw[:out_n.value]
        assert stream.endswith(b"\x03\x00")
        assert zlib.decompressobj(-12).decompress(stream) == data        # window 2^12 respected
        assert crc.value == zlib.crc32(data) and adler.value == zlib.adler32(data)
        assert cnt.value == (len(data) + BLOCK - 1) // BLOCK
    finally:
        m.close()


def test_two_real_devices_if_present(nat, single):
    if nat.lib().b2d_device_count() < 2:
        pytest.skip("one GPU on this box (the sharding itself is covered above with two contexts on device 0)")
    data = synthdata.make("mixed", 64 * MIB, 45).tobytes()
    want, want_crc, want_sizes = single_call(nat, single, data, 6)
    m = nat.MultiContext([0, 1], 8 * MIB)
    try:
        got, crc, sizes = multi_call(nat, m, data, len(data), 6)
    finally:
        m.close()
    assert got == want and crc == want_crc and sizes == want_sizes
