"""GPU parity tests of the block-parallel inflate (b2d_inflate_device / b2d_inflate_host):
streams written by this library's device-resident compressor, decoded with their block index, must give back
the input byte for byte — and the very same streams must decode to the same bytes with the checkers (oracle
inflate restated from inflate.c / inftrees.c, stdlib zlib).  All calls go through the C ABI."""
import ctypes
import zlib

import numpy as np
import pytest

import synthdata
from helpers import orc_inflate

pytestmark = pytest.mark.gpu
BLOCK = 131072


@pytest.fixture(scope="module")
def nat():
    from zlib_ng_b200 import _native
    return _native


@pytest.fixture(scope="module")
def ctx(nat):
    c = nat.Context(0, 8 << 20)      # 8 MiB batches: inputs above that exercise the index across batches
    yield c
    c.close()


def deflate_indexed(nat, ctx, data: bytes, level: int, zdict: bytes = b""):
    """device-resident compress of [zdict | data] -> (stream bytes, block offsets, device buffers)"""
    L = nat.lib()
    n, dn = len(data), len(zdict)
    d_buf = ctx.device_alloc(dn + n + 64 + 16)
    both = zdict + data
    if both:
        ctx.h2d(d_buf, both, len(both))
    cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
    d_out = ctx.device_alloc(cap + 16)
    out_n, crc = ctypes.c_uint64(0), ctypes.c_uint32(0)
    nat.check(L.b2d_deflate_device(ctx.handle, level, ctypes.c_void_p(d_buf), dn, n, ctypes.c_void_p(d_out), cap,
                                   ctypes.byref(out_n), ctypes.byref(crc), None))
    cnt = ctypes.c_uint64(0)
    nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
    offs = (ctypes.c_uint64 * max(cnt.value, 1))()
    nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
    stream = ctypes.create_string_buffer(max(out_n.value, 1))
    ctx.d2h(stream, d_out, out_n.value)
    return stream.raw[:out_n.value], list(offs)[:cnt.value], d_buf, d_out, crc.value


def inflate_device(nat, ctx, d_stream, stream_n, offs, d_dict, dict_n, raw_n):
    L = nat.lib()
    nb = (raw_n + BLOCK - 1) // BLOCK
    d_raw = ctx.device_alloc(raw_n + 64)
    arr = (ctypes.c_uint64 * max(nb, 1))(*offs[:nb])
    ms = ctypes.c_float(0)
    try:
        nat.check(L.b2d_inflate_device(ctx.handle, ctypes.c_void_p(d_stream), stream_n, arr, nb,
                                       ctypes.c_void_p(d_dict) if dict_n else None, dict_n, ctypes.c_void_p(d_raw), raw_n,
                                       ctypes.byref(ms)))
        back = ctypes.create_string_buffer(max(raw_n, 1))
        ctx.d2h(back, d_raw, raw_n)
    finally:
        ctx.device_free(d_raw)
    return back.raw[:raw_n]


CASES = [("fastq", 1_000_000, 5), ("mixed", 3_000_000, 6), ("fastq", 131072, 7), ("mixed", 131073, 8),
         ("fastq", 32769, 10), ("mixed", 1023, 11), ("mixed", 1, 12), ("mixed", 20_000_000, 13)]


@pytest.mark.parametrize("kind,n,seed", CASES)
@pytest.mark.parametrize("level", [0, 1, 6, 9, 6 | (4 << 8)])     # stored, dynamic, and Z_FIXED -> fixed-Huffman blocks
def test_round_trip_with_index(nat, ctx, kind, n, seed, level):
    if n > 4_000_000 and level not in (1, 6):
        pytest.skip("large case at two levels only")
    data = synthdata.make(kind, n, seed).tobytes()
    stream, offs, d_buf, d_out, crc = deflate_indexed(nat, ctx, data, level)
    try:
        assert len(offs) == (n + BLOCK - 1) // BLOCK + 1 and offs[0] == 0 and offs[-1] == len(stream)
        assert offs == sorted(offs)
        back = inflate_device(nat, ctx, d_out, len(stream), offs, 0, 0, n)
    finally:
        ctx.device_free(d_buf)
        ctx.device_free(d_out)
    assert back == data
    assert crc == zlib.crc32(data)
    # the checkers agree on what this stream holds
    assert zlib.decompressobj(-15).decompress(stream) == data
    if n <= 3_000_000:
        assert orc_inflate(stream, n, need_final=False) == data


@pytest.mark.parametrize("level", [1, 6])
def test_dictionary_and_markers_across_many_blocks(nat, ctx, level):
    # text that repeats with a period just under the window, runs of one byte, and a preset dictionary:
    # most of every block depends (transitively) on the 32 KiB before it
    rng = np.random.default_rng(3)
    period = rng.integers(32, 127, 30_000, dtype=np.uint8).tobytes()
    zdict = period[-20_000:]
    data = period * 20 + b"\x00" * 700_000 + period[:5000] * 40 + b"ab" * 300_000
    stream, offs, d_buf, d_out, _ = deflate_indexed(nat, ctx, data, level, zdict)
    try:
        back = inflate_device(nat, ctx, d_out, len(stream), offs, d_buf, len(zdict), len(data))
    finally:
        ctx.device_free(d_buf)
        ctx.device_free(d_out)
    assert back == data
    assert zlib.decompressobj(-15, zdict=zdict).decompress(stream) == data


def test_host_buffers_and_bad_index(nat, ctx):
    L = nat.lib()
    data = synthdata.make("fastq", 700_000, 21).tobytes()
    stream, offs, d_buf, d_out, _ = deflate_indexed(nat, ctx, data, 6)
    ctx.device_free(d_buf)
    ctx.device_free(d_out)
    nb = len(offs) - 1
    arr = (ctypes.c_uint64 * nb)(*offs[:nb])
    out = ctypes.create_string_buffer(len(data))
    crc = ctypes.c_uint32(0)
    nat.check(L.b2d_inflate_host(ctx.handle, stream, len(stream), arr, nb, None, 0, out, len(data), ctypes.byref(crc)))
    assert out.raw == data and crc.value == zlib.crc32(data)
    # an index that does not match the size, or points into the middle of a block, is refused / detected
    assert L.b2d_inflate_host(ctx.handle, stream, len(stream), arr, nb - 1, None, 0, out, len(data), None) == nat.B2D_STREAM_ERROR
    bad = (ctypes.c_uint64 * nb)(*([offs[0]] + [o + 3 for o in offs[1:nb]]))
    rc = L.b2d_inflate_host(ctx.handle, stream, len(stream), bad, nb, None, 0, out, len(data), None)
    assert rc == -3 or out.raw != data
    # empty
    nat.check(L.b2d_inflate_host(ctx.handle, None, 0, None, 0, None, 0, None, 0, None))


# ---- the user-facing side: gzip_ng_threaded.open(..., "rb") of files written by gzip_ng_threaded.open(..., "wb")
@pytest.fixture(scope="module")
def gzt():
    from zlib_ng_b200 import gzip_ng_threaded
    return gzip_ng_threaded


def test_reader_uses_the_block_index(gzt, tmp_path, monkeypatch):
    import gzip
    data = synthdata.make("mixed", 9_000_000, 31).tobytes()
    path = tmp_path / "indexed.gz"
    with gzt.open(path, "wb", compresslevel=6, threads=4) as f:
        for off in range(0, len(data), 1_000_003):
            f.write(data[off:off + 1_000_003])
    blob = path.read_bytes()
    assert gzip.decompress(blob) == data                   # index members are empty gzip members to everyone else
    monkeypatch.setattr(gzt, "_READ_WINDOW_BLOCKS", 8)      # several GPU calls, dictionary carried between them
    with gzt.open(path, "rb") as f:
        assert f.raw._gpu is not None
        back = f.read()
    assert back == data
    with gzt.open(path, "rt", encoding="latin-1") as f:
        assert f.read(100) == data[:100].decode("latin-1")
    # a flipped bit in the compressed data is caught (decode error or CRC mismatch), like gzip does
    bad = bytearray(blob)
    bad[len(blob) // 3] ^= 0x10
    (tmp_path / "bad.gz").write_bytes(bytes(bad))
    with pytest.raises((gzip.BadGzipFile, EOFError, OSError)):
        with gzt.open(tmp_path / "bad.gz", "rb") as f:
            f.read()


def test_reader_without_index_and_small_files(gzt, tmp_path):
    import gzip
    data = synthdata.make("fastq", 300_000, 32).tobytes()
    (tmp_path / "plain.gz").write_bytes(gzip.compress(data))
    with gzt.open(tmp_path / "plain.gz", "rb") as f:
        assert f.raw._gpu is None
        assert f.read() == data
    for n in (1, 131072, 131073):
        with gzt.open(tmp_path / "s.gz", "wb", threads=2) as f:
            f.write(data[:n])
        with gzt.open(tmp_path / "s.gz", "rb") as f:
            assert f.raw._gpu is not None
            assert f.read() == data[:n]
    # flush() starts a new member: such files carry no index and are read the ordinary way
    with gzt.open(tmp_path / "two.gz", "wb", threads=2) as f:
        f.write(data)
        f.flush()
        f.write(data)
    with gzt.open(tmp_path / "two.gz", "rb") as f:
        assert f.raw._gpu is None
        assert f.read() == data + data
    # empty file
    with gzt.open(tmp_path / "e.gz", "wb", threads=2) as f:
        pass
    with gzt.open(tmp_path / "e.gz", "rb") as f:
        assert f.read() == b""


@pytest.mark.timeout(60)
def test_reader_can_be_closed_early(gzt, tmp_path, monkeypatch):
    import threading
    data = synthdata.make("fastq", 6_000_000, 33).tobytes()
    path = tmp_path / "early.gz"
    with gzt.open(path, "wb", threads=2) as f:
        f.write(data)
    monkeypatch.setattr(gzt, "_READ_WINDOW_BLOCKS", 4)
    before = threading.active_count()
    f = gzt.open(path, "rb")
    assert f.read(1000) == data[:1000]
    f.close()                                   # the decoder threads are one window ahead: they must stop, not hang
    assert f.closed and threading.active_count() <= before
    with gzt.open(path, "rb") as f:             # and the pooled buffers serve the next reader
        assert f.read() == data


@pytest.mark.timeout(120)
def test_writers_and_readers_sharing_the_contexts(gzt, tmp_path, monkeypatch):
    """Several files written, then read, at the same time in one process: the GPU contexts are shared, so every
    block index has to come from its own compress call and every host inflate owns the context's buffers."""
    import threading
    monkeypatch.setattr(gzt, "_READ_WINDOW_BLOCKS", 16)
    datas = [synthdata.make(kind, 5_000_000 + 70_001 * i, 40 + i).tobytes() for i, kind in enumerate(("fastq", "mixed", "fastq"))]
    errors = []

    def write(i):
        try:
            with gzt.open(tmp_path / f"w{i}.gz", "wb", threads=2) as f:
                for off in range(0, len(datas[i]), 300_000):
                    f.write(datas[i][off:off + 300_000])
        except BaseException as e:
            errors.append(e)

    def read(i):
        try:
            with gzt.open(tmp_path / f"w{i}.gz", "rb") as f:
                assert f.raw._gpu is not None
                assert f.read() == datas[i]
        except BaseException as e:
            errors.append(e)

    for fn in (write, read):
        ts = [threading.Thread(target=fn, args=(i,)) for i in range(len(datas))]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errors, errors


def test_reader_interface_on_indexed_files(gzt):
    """The reference's reader tests (tests/test_gzip_ng_threaded.py:73-80,114-136,180-205) on the GPU path."""
    import io
    data = synthdata.make("fastq", 700_000, 50).tobytes()
    sink = io.BytesIO()
    with gzt.open(sink, "wb", threads=2) as f:
        f.write(data)
    assert not sink.closed                                  # streams passed in are not closed
    blob = sink.getvalue()
    stream = io.BytesIO(blob)
    with gzt.open(stream, "rb") as f:
        assert f.raw._gpu is not None
        assert f.readable() and not f.writable()
        assert f.read(10) == data[:10] and f.tell() == 10
        assert f.read() == data[10:]
        assert f.read() == b"" and f.tell() == len(data)
    assert not stream.closed
    r = gzt._ThreadedGzipReader(io.BytesIO(blob))
    r.close()
    assert r.closed
    r.close()                                               # double close is fine
    with pytest.raises(ValueError, match="closed"):
        r.readinto(bytearray(10))
    with gzt.open(io.BytesIO(blob[:-8]), "rb") as f:        # truncated: no locator any more, CPython's reader reports it
        with pytest.raises(EOFError):
            f.read()
    # an index that no longer fits the data (trailer CRC changed) is reported like a damaged gzip file
    import gzip
    import struct
    end = gzt._LOCATOR.unpack(gzt._parse_extra_member(blob[-gzt._LOCATOR_MEMBER:], 0)[0])[4]
    empty = 10 + 2 + 8
    pos = end - empty - 8
    bad = blob[:pos] + struct.pack("<I", struct.unpack_from("<I", blob, pos)[0] ^ 1) + blob[pos + 4:]
    with gzt.open(io.BytesIO(bad), "rb") as f:
        with pytest.raises(gzip.BadGzipFile, match="CRC"):
            f.read()


def test_gzip_ng_decompress_of_indexed_data(gzt):
    import gzip
    import io
    from zlib_ng_b200 import gzip_ng
    data = synthdata.make("mixed", 2_000_000, 51).tobytes()
    sink = io.BytesIO()
    with gzt.open(sink, "wb", threads=2) as f:
        f.write(data)
    blob = sink.getvalue()
    r = gzt._ThreadedGzipReader(io.BytesIO(blob))
    assert r._gpu is not None
    r.close()
    assert gzip_ng.decompress(blob) == data
    assert gzip_ng.decompress(gzip.compress(data[:1000])) == data[:1000]        # no index: CPython


def test_failed_block_never_leaves_stale_symbols_behind(nat):
    """the 16-bit symbol scratch is reused across calls.  A dictionary-primed inflate leaves markers
    below 32768 in it; a later call WITHOUT a dictionary whose block 0 fails part-way must not resolve those stale
    markers (it used to read dict[-32768 + k] from a null pointer: a sticky illegal address).  The context must
    report bad data and go on working."""
    L = nat.lib()
    c = nat.Context(0, 8 << 20)
    try:
        rng = np.random.default_rng(9)
        period = rng.integers(32, 127, 30_000, dtype=np.uint8).tobytes()
        zdict = period[-20_000:]
        data = period * 12                       # nearly every symbol of block 0 refers into the dictionary
        stream, offs, d_buf, d_out, _ = deflate_indexed(nat, c, data, 6, zdict)
        try:
            back = inflate_device(nat, c, d_out, len(stream), offs, d_buf, len(zdict), len(data))
            assert back == data
        finally:
            c.device_free(d_buf)
            c.device_free(d_out)
        # same size, no dictionary, block 0 damaged right after its first header
        plain = synthdata.make("mixed", len(data), 61).tobytes()
        stream2, offs2, d_buf2, d_out2, _ = deflate_indexed(nat, c, plain, 6)
        c.device_free(d_buf2)
        c.device_free(d_out2)
        nb = len(offs2) - 1
        arr = (ctypes.c_uint64 * nb)(*offs2[:nb])
        out = ctypes.create_string_buffer(len(plain))
        for cut in (40, 200, 5000):
            bad = bytearray(stream2)
            bad[cut:cut + 64] = b"\xff" * 64
            rc = L.b2d_inflate_host(c.handle, bytes(bad), len(bad), arr, nb, None, 0, out, len(plain), None)
            assert rc == -3 or out.raw != plain              # detected (or, at worst, different bytes: the CRC catches it)
        # the context still works: nothing sticky happened on the device
        nat.check(L.b2d_inflate_host(c.handle, stream2, len(stream2), arr, nb, None, 0, out, len(plain), None))
        assert out.raw == plain
    finally:
        c.close()


@pytest.mark.timeout(120)
@pytest.mark.parametrize("env", [{"CUDA_LAUNCH_BLOCKING": "1"}, {"B2D_INF_SERIAL": "1"}, {}])
def test_inflate_does_not_depend_on_two_kernels_running_at_once(tmp_path, env):
    """the pass over the tails used to be launched first and to wait for a decode kernel that, when
    launches are serialised (CUDA_LAUNCH_BLOCKING=1, profilers), could not start: 20 s of spinning, then a false
    Z_DATA_ERROR.  Now the decode goes first and waits for nobody; a valid file reads correctly, and fast, however
    kernels are scheduled."""
    import os
    import subprocess
    import sys
    import time
    code = r'''
import sys, time
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import synthdata
from zlib_ng_b200 import gzip_ng_threaded as gzt
data = synthdata.make("mixed", 6_000_000, 71).tobytes()
with gzt.open(sys.argv[3], "wb", threads=2) as f:
    f.write(data)
t0 = time.perf_counter()
with gzt.open(sys.argv[3], "rb") as f:
    assert f.raw._gpu is not None
    back = f.read()
dt = time.perf_counter() - t0
assert back == data
print("READ_SECONDS", dt)
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    e = dict(os.environ)
    e.update(env)
    t0 = time.perf_counter()
    p = subprocess.run([sys.executable, "-c", code, root, os.path.join(root, "python-zlib-ng_b200"), str(tmp_path / "f.gz")],
                       env=e, capture_output=True, text=True, timeout=100)
    assert p.returncode == 0, p.stderr[-2000:]
    secs = float(p.stdout.split("READ_SECONDS")[1].split()[0])
    assert secs < 2.0, f"reading took {secs:.2f} s"
    del t0


# ---- BGZF (blocked gzip, htslib / bgzip): members carry their size in a "BC" extra subfield; the reference reads its
# fixture tests/data/test.fastq.bgzip.gz with one inflate thread (tests/test_gzip_ng.py:466-473)
def make_bgzf(data: bytes, blk: int = 65280, level: int = 6, eof_marker: bool = True) -> bytes:
    import struct
    out = []
    pieces = [data[o:o + blk] for o in range(0, len(data), blk)] + ([b""] if eof_marker else [])
    for piece in pieces:
        c = zlib.compressobj(level, zlib.DEFLATED, -15)
        body = c.compress(piece) + c.flush()
        out.append(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", len(body) + 25) + body
                   + struct.pack("<II", zlib.crc32(piece), len(piece)))
    return b"".join(out)


def test_bgzf_files_are_inflated_on_the_gpu(gzt, tmp_path, monkeypatch):
    import gzip
    import io
    data = synthdata.make("fastq", 9_000_000, 81).tobytes() + synthdata.make("mixed", 3_000_000, 82).tobytes()
    blob = make_bgzf(data)
    assert gzip.decompress(blob) == data                            # an ordinary multi-member gzip file to everyone else
    path = tmp_path / "x.bgzip.gz"
    path.write_bytes(blob)
    for window in (1 << 20, 32 << 20):                              # a small window: members carried over between calls
        monkeypatch.setattr(gzt, "_BGZF_WINDOW", window)
        with gzt.open(path, "rb") as f:
            assert f.raw._gpu is not None and f.raw._gpu["kind"] == "bgzf"
            assert f.read() == data
        with gzt.open(path, "rb") as f:
            assert f.read(1000) == data[:1000] and f.tell() == 1000
            assert f.read(5_000_000) == data[1000:5_001_000]
    with gzt.open(path, "rt", encoding="latin-1") as f:
        assert f.readline() == data[:data.index(b"\n") + 1].decode("latin-1")
    # stored members (level 0) and members of a few bytes
    small = make_bgzf(data[:200_000], blk=1000, level=0)
    with gzt.open(io.BytesIO(small), "rb") as f:
        assert f.raw._gpu["kind"] == "bgzf" and f.read() == data[:200_000]
    # a flipped bit inside a member: decode error or CRC mismatch, reported like a damaged gzip file
    bad = bytearray(blob)
    bad[len(blob) // 2] ^= 0x04
    with gzt.open(io.BytesIO(bytes(bad)), "rb") as f:
        with pytest.raises((gzip.BadGzipFile, EOFError, OSError, zlib.error)):
            f.read()
    # an ordinary gzip member appended to a BGZF file: the rest is read the ordinary way
    mixed = make_bgzf(data[:3_000_000], eof_marker=False) + gzip.compress(data[3_000_000:4_000_000])
    with gzt.open(io.BytesIO(mixed), "rb") as f:
        assert f.raw._gpu["kind"] == "bgzf" and f.read() == data[:4_000_000]
    # a file cut off inside a member: EOFError, as gzip reports it
    with gzt.open(io.BytesIO(blob[:len(blob) // 2]), "rb") as f:
        with pytest.raises(EOFError):
            f.read()


def test_inflate_members_through_the_c_abi(nat, ctx):
    L = nat.lib()
    data = synthdata.make("mixed", 2_000_000, 83).tobytes()
    blob = make_bgzf(data)
    n = len(blob)
    m = n // 28 + 16
    po, pl = (ctypes.c_uint64 * m)(), (ctypes.c_uint32 * m)()
    rl, cr = (ctypes.c_uint32 * m)(), (ctypes.c_uint32 * m)()
    cnt, used = ctypes.c_uint64(0), ctypes.c_uint64(0)
    nat.check(L.b2d_bgzf_scan(blob, n, m, po, pl, rl, cr, ctypes.byref(cnt), ctypes.byref(used)))
    assert used.value == n and sum(rl[:cnt.value]) == len(data)
    out = ctypes.create_string_buffer(len(data))
    out_n = ctypes.c_uint64(0)
    nat.check(L.b2d_inflate_members_host(ctx.handle, blob, n, po, pl, rl, cr, cnt.value, out, len(data), ctypes.byref(out_n)))
    assert out_n.value == len(data) and out.raw == data
    # wrong CRC for one member: -3, like a damaged member
    cr[3] ^= 1
    assert L.b2d_inflate_members_host(ctx.handle, blob, n, po, pl, rl, cr, cnt.value, out, len(data), ctypes.byref(out_n)) == -3
    assert "CRC" in nat.last_error()
    cr[3] ^= 1
    # output buffer too small, and a member that claims more than it holds
    assert L.b2d_inflate_members_host(ctx.handle, blob, n, po, pl, rl, cr, cnt.value, out, 1000, ctypes.byref(out_n)) == nat.B2D_BUF_ERROR
    rl[2] -= 1
    assert L.b2d_inflate_members_host(ctx.handle, blob, n, po, pl, rl, None, cnt.value, out, len(data), ctypes.byref(out_n)) == -3
    rl[2] += 1
    nat.check(L.b2d_inflate_members_host(ctx.handle, blob, n, po, pl, rl, cr, cnt.value, out, len(data), ctypes.byref(out_n)))
    assert out.raw == data


def test_host_inflate_over_many_chunks(nat):
    """b2d_inflate_host pipelines chunks of 512 blocks (H2D / decode / resolve / D2H side by side, the chain over the
    block tails running through all of them): 1300 blocks = 3 chunks, with a dictionary, CRC continued on the device."""
    L = nat.lib()
    c = nat.Context(0, 64 << 20)
    try:
        n = 1300 * BLOCK - 4321
        base = synthdata.make("mixed", 24_000_000, 84).tobytes()
        data = (base * (n // len(base) + 1))[:n]
        zdict = base[-30_000:]
        stream, offs, d_buf, d_out, _ = deflate_indexed(nat, c, data, 6, zdict)
        c.device_free(d_buf)
        c.device_free(d_out)
        nb = len(offs) - 1
        arr = (ctypes.c_uint64 * nb)(*offs[:nb])
        out = ctypes.create_string_buffer(n)
        for rep in range(2):
            crc = ctypes.c_uint32(zlib.crc32(b"before"))
            nat.check(L.b2d_inflate_host(c.handle, stream, len(stream), arr, nb, zdict, len(zdict), out, n, ctypes.byref(crc)))
            assert out.raw == data
            assert crc.value == zlib.crc32(data, zlib.crc32(b"before"))
    finally:
        c.close()


# ---- one warp per 32 KiB sub-block: the writer's bit offsets of the sub-blocks (b2d_subblock_bits, index kind 2)
def sub_bits_of_last_call(nat, ctx, nb):
    L = nat.lib()
    cnt = ctypes.c_uint64(0)
    nat.check(L.b2d_subblock_bits(ctx.handle, None, 0, ctypes.byref(cnt)))
    assert cnt.value == 4 * nb
    bits = (ctypes.c_uint32 * max(cnt.value, 1))()
    nat.check(L.b2d_subblock_bits(ctx.handle, bits, cnt.value, ctypes.byref(cnt)))
    return bits


@pytest.mark.parametrize("kind,n,seed", [("fastq", 1_000_000, 5), ("mixed", 3_000_000, 6), ("mixed", 131073, 8), ("fastq", 32769, 10),
                                         ("mixed", 1023, 11), ("mixed", 20_000_000, 13), ("fastq", 98_304, 14), ("fastq", 98_305, 15)])
@pytest.mark.parametrize("level", [0, 1, 6, 6 | (4 << 8)])
def test_round_trip_one_warp_per_sub_block(nat, ctx, kind, n, seed, level):
    if n > 4_000_000 and level != 6:
        pytest.skip("large case at one level only")
    L = nat.lib()
    data = synthdata.make(kind, n, seed).tobytes()
    zdict = synthdata.make("fastq", 20_000, seed + 100).tobytes() if seed % 2 else b""
    stream, offs, d_buf, d_out, _ = deflate_indexed(nat, ctx, data, level, zdict)
    nb = len(offs) - 1
    bits = sub_bits_of_last_call(nat, ctx, nb)
    try:
        for b in range(nb):                                   # first sub-block at bit 0, the others strictly after
            row = list(bits[4 * b:4 * b + 4])
            have = -(-min(BLOCK, n - b * BLOCK) // 32768)
            assert row[0] == 0 and all(row[i] > row[i - 1] for i in range(1, have)) and all(v == 0 for v in row[have:])
        arr = (ctypes.c_uint64 * nb)(*offs[:nb])
        d_raw = ctx.device_alloc(n + 64)
        try:
            nat.check(L.b2d_inflate_device_sub(ctx.handle, ctypes.c_void_p(d_out), len(stream), arr, bits, nb,
                                               ctypes.c_void_p(d_buf) if zdict else None, len(zdict), ctypes.c_void_p(d_raw), n, None))
            back = ctypes.create_string_buffer(n)
            ctx.d2h(back, d_raw, n)
        finally:
            ctx.device_free(d_raw)
        assert back.raw == data
        # host form, CRC continued on the device
        out = ctypes.create_string_buffer(n)
        crc = ctypes.c_uint32(0)
        nat.check(L.b2d_inflate_host_sub(ctx.handle, stream, len(stream), arr, bits, nb, zdict if zdict else None, len(zdict), out, n,
                                         ctypes.byref(crc)))
        assert out.raw == data and crc.value == zlib.crc32(data)
        if nb >= 1 and n > 40_000 and (level & 0xff) != 0:      # (stored blocks begin with padding: three bits more still decode)
            # a sub-block offset that points into the middle of a code: detected, or at worst different bytes
            bad = (ctypes.c_uint32 * (4 * nb))(*bits)
            bad[1] += 3
            rc = L.b2d_inflate_host_sub(ctx.handle, stream, len(stream), arr, bad, nb, zdict if zdict else None, len(zdict), out, n, None)
            assert rc == -3 or out.raw != data
            bad[1] = 9 * (BLOCK + 1024)                       # out of range: refused before anything runs
            assert L.b2d_inflate_host_sub(ctx.handle, stream, len(stream), arr, bad, nb, None, 0, out, n, None) == nat.B2D_STREAM_ERROR
    finally:
        ctx.device_free(d_buf)
        ctx.device_free(d_out)


def test_marker_chains_through_sub_blocks(nat, ctx):
    """runs and a period just under the window: nearly every symbol of every sub-block refers to the sub-block before
    it, through all four sub-blocks of a block and on into the next block"""
    L = nat.lib()
    rng = np.random.default_rng(4)
    period = rng.integers(32, 127, 30_000, dtype=np.uint8).tobytes()
    zdict = period[-20_000:]
    data = period * 20 + b"\x00" * 700_000 + period[:5000] * 40 + b"ab" * 300_000
    for level in (1, 6):
        stream, offs, d_buf, d_out, _ = deflate_indexed(nat, ctx, data, level, zdict)
        ctx.device_free(d_buf)
        ctx.device_free(d_out)
        nb = len(offs) - 1
        bits = sub_bits_of_last_call(nat, ctx, nb)
        arr = (ctypes.c_uint64 * nb)(*offs[:nb])
        out = ctypes.create_string_buffer(len(data))
        nat.check(L.b2d_inflate_host_sub(ctx.handle, stream, len(stream), arr, bits, nb, zdict, len(zdict), out, len(data), None))
        assert out.raw == data


def test_written_files_carry_sub_block_offsets(gzt, tmp_path, monkeypatch):
    import gzip
    data = synthdata.make("mixed", 9_000_000, 91).tobytes()
    p = tmp_path / "v2.gz"
    with gzt.open(p, "wb", threads=4) as f:
        f.write(data)
    assert gzip.decompress(p.read_bytes()) == data
    monkeypatch.setattr(gzt, "_READ_WINDOW_BLOCKS", 16)
    with gzt.open(p, "rb") as f:
        assert f.raw._gpu is not None and f.raw._gpu["subs"] is not None and len(f.raw._gpu["subs"]) == 4 * len(f.raw._gpu["sizes"])
        assert f.read() == data
    # files with the older index (block sizes only) are still read on the GPU, one warp per block
    monkeypatch.setattr(gzt, "_WRITE_SUBS", False)
    q = tmp_path / "v1.gz"
    with gzt.open(q, "wb", threads=4) as f:
        f.write(data)
    with gzt.open(q, "rb") as f:
        assert f.raw._gpu is not None and f.raw._gpu["subs"] is None
        assert f.read() == data


# ---- the host call in two halves: b2d_inflate_host_decode (no dictionary, no waiting) / b2d_inflate_host_resolve
def test_two_windows_decoded_ahead_and_resolved_in_order(nat):
    """What the reader does: the stream is cut into two windows at a block boundary; both decodes are queued (on two
    contexts) before the first window is resolved, the second one against the last 32 KiB of the first.  The CRC runs on.
    Then the error behaviour of the pair of calls."""
    L = nat.lib()
    a, b = nat.Context(0, 1 << 20), nat.Context(0, 1 << 20)
    comp = nat.Context(0, 64 << 20)
    try:
        n = 700 * BLOCK + 12345
        base = synthdata.make("mixed", 16_000_000, 91).tobytes()
        data = (base * (n // len(base) + 1))[:n]
        zdict = base[-20_000:]
        stream, offs, d_buf, d_out, _ = deflate_indexed(nat, comp, data, 6, zdict)
        bits = sub_bits_of_last_call(nat, comp, len(offs) - 1)
        comp.device_free(d_buf)
        comp.device_free(d_out)
        nb = len(offs) - 1
        cut = 300                                      # blocks in the first window
        w1 = stream[:offs[cut]]
        w2 = stream[offs[cut]:]
        o1 = (ctypes.c_uint64 * cut)(*offs[:cut])
        o2 = (ctypes.c_uint64 * (nb - cut))(*[o - offs[cut] for o in offs[cut:nb]])
        s1 = (ctypes.c_uint32 * (4 * cut))(*bits[:4 * cut])
        s2 = (ctypes.c_uint32 * (4 * (nb - cut)))(*bits[4 * cut:4 * nb])
        n1, n2 = cut * BLOCK, n - cut * BLOCK
        out1, out2 = ctypes.create_string_buffer(n1), ctypes.create_string_buffer(n2)
        for subs1, subs2 in ((s1, s2), (None, None)):
            nat.check(L.b2d_inflate_host_decode(a.handle, w1, len(w1), o1, subs1, cut, len(zdict), n1))
            nat.check(L.b2d_inflate_host_decode(b.handle, w2, len(w2), o2, subs2, nb - cut, 32768, n2))
            crc = ctypes.c_uint32(0)
            nat.check(L.b2d_inflate_host_resolve(a.handle, zdict, len(zdict), out1, n1, ctypes.byref(crc)))
            assert out1.raw == data[:n1]
            nat.check(L.b2d_inflate_host_resolve(b.handle, out1.raw[-32768:], 32768, out2, n2, ctypes.byref(crc)))
            assert out2.raw == data[n1:]
            assert crc.value == zlib.crc32(data)
        # a second decode before the resolve, a resolve without a decode, a dictionary of another length: refused
        nat.check(L.b2d_inflate_host_decode(a.handle, w1, len(w1), o1, s1, cut, len(zdict), n1))
        assert L.b2d_inflate_host_decode(a.handle, w1, len(w1), o1, s1, cut, len(zdict), n1) == nat.B2D_STREAM_ERROR
        assert L.b2d_inflate_host_resolve(a.handle, zdict, len(zdict) - 1, out1, n1, None) == nat.B2D_STREAM_ERROR
        assert L.b2d_inflate_host_resolve(a.handle, zdict, len(zdict), out1, n1, None) == nat.B2D_STREAM_ERROR   # given up by the mismatch
        # giving a queued decode up leaves the context usable
        nat.check(L.b2d_inflate_host_decode(a.handle, w1, len(w1), o1, s1, cut, len(zdict), n1))
        nat.check(L.b2d_inflate_host_resolve(a.handle, None, 0, None, 0, None))
        nat.check(L.b2d_inflate_host_sub(a.handle, w1, len(w1), o1, s1, cut, zdict, len(zdict), out1, n1, None))
        assert out1.raw == data[:n1]
    finally:
        a.close()
        b.close()
        comp.close()
