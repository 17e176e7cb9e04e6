"""GPU tests of the reference-facing Python API (zlib_ng_b200.zlib_ng,
gzip_ng_threaded, gzip_ng).  They follow the reference's own tests
(tests/test_gzip_ng_threaded.py, tests/test_compat.py and the checksum /
compress cases of tests/test_zlib_compliance.py) with stdlib zlib / gzip as the
cross-check decoder, as those tests do."""
import copy
import gzip
import io
import itertools
import os
import pickle
import subprocess
import sys
import zlib

import numpy as np
import pytest

import synthdata

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def zng():
    from zlib_ng_b200 import zlib_ng
    return zlib_ng


@pytest.fixture(scope="module")
def gzt():
    from zlib_ng_b200 import gzip_ng_threaded
    return gzip_ng_threaded


@pytest.fixture(scope="module")
def fastq():
    return synthdata.make("fastq", 3_578_369, 77).tobytes()


# ---------------------------------------------------------------- checksums (test_zlib_compliance.py:76-127)
def test_checksum_known_answers(zng):
    assert zng.crc32(b"") == zng.crc32(b"", 0) == 0
    assert zng.crc32(b"", 1) == 1 and zng.crc32(b"", 432) == 432
    assert zng.adler32(b"") == zng.adler32(b"", 1) == 1
    assert zng.adler32(b"", 0) == 0 and zng.adler32(b"", 432) == 432
    assert zng.crc32(b"penguin", 0) == 0x0e5c1a120
    assert zng.crc32(b"penguin", 1) == 0x43b6aa94
    assert zng.adler32(b"penguin", 0) == 0x0bcf02f6
    assert zng.adler32(b"penguin", 1) == 0x0bd602f7
    assert zng.crc32(b"abcdefghijklmnop") == 2486878355
    assert zng.crc32(b"spam") == 1138425661
    foo = b"abcdefghijklmnop"
    assert zng.adler32(foo + foo) == 3573550353
    assert zng.adler32(b"spam") == 72286642


def test_crc32_combine(zng):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, 5000, dtype=np.uint8).tobytes()
    b = rng.integers(0, 256, 77777, dtype=np.uint8).tobytes()
    assert zng.crc32_combine(zng.crc32(a), zng.crc32(b), len(b)) == zlib.crc32(a + b)
    assert zng.crc32_combine(0, zng.crc32(b), len(b)) == zlib.crc32(b)
    with pytest.raises(TypeError):
        zng.crc32_combine(1.5, 2, 3)
    with pytest.raises(OverflowError):
        zng.crc32_combine(-1, 2, 3)


DATA_SIZES = [8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072, 262144, 524288]
SEEDS = [0, 1, 2, 3, 65521, 65522, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF, -1, -65521, 2 ** 64 + 5, -(2 ** 40)]


@pytest.mark.parametrize("size", DATA_SIZES)
def test_checksums_vs_stdlib_with_wrapping_start_values(zng, size):
    # tests/test_compat.py:26-32,66-77: start values wrap mod 2^32 (PyLong_AsUnsignedLongMask)
    data = synthdata.make("mixed", size, size).tobytes()
    for s in SEEDS:
        assert zng.crc32(data, s) == zlib.crc32(data, s & 0xFFFFFFFF)
        assert zng.adler32(data, s) == zlib.adler32(data, s & 0xFFFFFFFF)


def test_checksum_buffer_types(zng):
    data = synthdata.make("fastq", 100_000, 1)
    want = zlib.crc32(data.tobytes())
    for obj in (data, data.tobytes(), bytearray(data.tobytes()), memoryview(data.tobytes()), memoryview(data)):
        assert zng.crc32(obj) == want
    with pytest.raises(TypeError):
        zng.crc32("a string")


# ---------------------------------------------------------------- compress (test_compat.py:80-128)
WBITS = list(range(9, 16)) + list(range(-15, -8)) + list(range(25, 32))


@pytest.mark.parametrize("level", range(-1, 10))
def test_compress_levels_round_trip(zng, fastq, level):
    data = fastq[:400_000]
    for wbits in (15, -15, 31):
        out = zng.compress(data, level=level, wbits=wbits)
        assert zlib.decompress(out, wbits=wbits) == data


@pytest.mark.parametrize("wbits", WBITS)
def test_compress_wbits_round_trip(zng, fastq, wbits):
    data = fastq[1_000_000:1_300_000]
    out = zng.compress(data, level=6, wbits=wbits)
    assert zlib.decompress(out, wbits=wbits) == data      # decoder limited to the same window


def test_compress_ratio_is_in_range(zng, fastq):
    # stdlib zlib (1.3) level 6 is the stand-in decoder-side yardstick on the box; the reference's
    # own sizes are pinned in tests/golden (test_ratio_vs_reference_golden)
    ours = len(zng.compress(fastq, 6, -15))
    theirs = len(zlib.compress(fastq, 6, wbits=-15))
    assert ours <= 1.03 * theirs


def test_compress_bad_arguments(zng):
    with pytest.raises(zng.error):
        zng.compress(b"ERROR", 10)
    with pytest.raises(zng.error):
        zng.compress(b"ERROR", 6, 42)
    with pytest.raises(TypeError):
        zng.compress()
    with pytest.raises(ValueError):
        zng.compressobj(1, zng.DEFLATED, 0)
    with pytest.raises(ValueError):
        zng.compressobj(6, 9)
    with pytest.raises(ValueError):
        zng.compressobj(6, zng.DEFLATED, 15, 10)


def test_gzip_and_zlib_headers(zng):
    data = b"headers" * 1000
    z = zng.compress(data, 6)
    assert z[:2] == zlib.compress(data, 6)[:2] == b"\x78\x9c"
    assert zng.compress(data, 1)[:2] == b"\x78\x01" and zng.compress(data, 9)[:2] == b"\x78\xda"
    g = zng.compress(data, 9, 31)
    assert g[:4] == b"\x1f\x8b\x08\x00" and g[8] == 2 and g[9] == 3
    assert zng.compress(data, 1, 31)[8] == 4
    assert gzip.decompress(g) == data


# ---------------------------------------------------------------- compressobj (test_zlib_compliance.py:226-306,503-608,711-789)
def test_pair_compress_equals_compressobj(zng, fastq):
    for data in (fastq[:5_000_000], b"", b"x", fastq[:131072], fastq[:131073]):
        for level, wbits in ((6, 15), (1, -15), (9, 31)):
            one = zng.compress(data, level, wbits)
            co = zng.compressobj(level, zng.DEFLATED, wbits)
            two = co.compress(data) + co.flush()
            assert one == two
            # and sliced writes produce the same stream again
            co = zng.compressobj(level, zng.DEFLATED, wbits)
            parts = [co.compress(data[i:i + 70_001]) for i in range(0, len(data), 70_001)]
            assert b"".join(parts) + co.flush() == one


def test_compressobj_streams_large_input_incrementally(zng, monkeypatch):
    monkeypatch.setattr(zng, "_STREAM_THRESHOLD", 4 << 20)
    monkeypatch.setattr(zng, "_PIN_BYTES", 2 << 20)      # the page-locked gather buffer long streams switch to
    data = synthdata.make("mixed", 12_000_000, 4).tobytes()
    co = zng.compressobj(6)
    pieces = [co.compress(data[i:i + 1_000_000]) for i in range(0, len(data), 1_000_000)]
    assert sum(1 for p in pieces if p) >= 2          # output appears before flush
    out = b"".join(pieces) + co.flush()
    assert zlib.decompress(out) == data
    assert out == zng.compress(data, 6)              # however the writes were sliced
    # one write larger than the threshold goes to the GPU straight from the caller's buffer
    co = zng.compressobj(6)
    first = co.compress(memoryview(data)[:9_000_001])
    assert first
    assert first + co.compress(data[9_000_001:]) + co.flush() == out
    with pytest.raises(zng.error):
        co.compress(b"more")


def test_compressobj_strategies_change_the_stream(zng):
    """strategy is honoured, not just accepted (deflate_huff.c / deflate_rle.c / trees.c:677)"""
    data = b"".join(bytes([i % 251]) * 900 for i in range(400)) + b"abcdefgh" * 20_000
    size = {}
    for strat in (zng.Z_DEFAULT_STRATEGY, zng.Z_FILTERED, zng.Z_HUFFMAN_ONLY, zng.Z_RLE, zng.Z_FIXED):
        co = zng.compressobj(6, zng.DEFLATED, -15, 8, strat)
        out = co.compress(data) + co.flush()
        assert zlib.decompress(out, wbits=-15) == data
        size[strat] = len(out)
        if strat == zng.Z_FIXED:
            assert (out[0] >> 1) & 3 in (0, 1)
    assert size[zng.Z_HUFFMAN_ONLY] > 10 * size[zng.Z_DEFAULT_STRATEGY]
    assert size[zng.Z_DEFAULT_STRATEGY] < size[zng.Z_RLE] < size[zng.Z_HUFFMAN_ONLY]


@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_flush_modes(zng, level):
    data = synthdata.make("fastq", 300_000, 8).tobytes()
    for sync in (zng.Z_NO_FLUSH, zng.Z_PARTIAL_FLUSH, zng.Z_SYNC_FLUSH, zng.Z_FULL_FLUSH, zng.Z_BLOCK):
        co = zng.compressobj(level)
        a = co.compress(data[:100_000])
        b = co.flush(sync)
        c = co.compress(data[100_000:])
        d = co.flush()
        assert zlib.decompress(a + b + c + d) == data
        if sync in (zng.Z_SYNC_FLUSH, zng.Z_FULL_FLUSH):
            # everything written so far is decodable after a sync flush
            assert zlib.decompressobj().decompress(a + b) == data[:100_000]
            assert (a + b).endswith(b"\x00\x00\xff\xff")


def test_zdict(zng):
    zdict = b"".join([b"line %d of the dictionary\n" % i for i in range(2000)])
    data = b"".join([b"line %d of the dictionary\n" % i for i in range(1500, 2500)])
    for wbits in (15, -15):
        co = zng.compressobj(6, zng.DEFLATED, wbits, zdict=zdict)
        out = co.compress(data) + co.flush()
        do = zlib.decompressobj(wbits, zdict=zdict)
        assert do.decompress(out) + do.flush() == data
        plain = zng.compress(data, 6, wbits)
        assert len(out) < len(plain)
    with pytest.raises(TypeError):
        zng.compressobj(zdict="not bytes")


def test_compressobj_copy_and_pickle(zng):
    d0, d1 = synthdata.make("fastq", 200_000, 1).tobytes(), synthdata.make("mixed", 200_000, 2).tobytes()
    for fn in (lambda c: c.copy(), copy.copy, copy.deepcopy):
        c0 = zng.compressobj(6)
        bufs0 = [c0.compress(d0)]
        c1 = fn(c0)
        bufs1 = list(bufs0)
        bufs0 += [c0.compress(d0), c0.flush()]
        bufs1 += [c1.compress(d1), c1.flush()]
        assert zlib.decompress(b"".join(bufs0)) == d0 + d0
        assert zlib.decompress(b"".join(bufs1)) == d0 + d1
    c = zng.compressobj()
    c.flush()
    with pytest.raises(ValueError):
        c.copy()
    for proto in range(pickle.HIGHEST_PROTOCOL + 1):
        with pytest.raises((TypeError, pickle.PicklingError)):
            pickle.dumps(zng.compressobj(), proto)


# ---------------------------------------------------------------- _ParallelCompress
def test_parallel_compress(zng, fastq):
    bs = 128 * 1024
    pc = zng._ParallelCompress(bs + max(bs // 10, 500), 1)
    prev, out, crc, total = b"", [], 0, 0
    for i in range(0, 1_000_000, bs):
        blk = fastq[i:i + bs]
        comp, c = pc.compress_and_crc(blk, prev[-32768:])
        assert comp.endswith(b"\x00\x00\xff\xff") and c == zlib.crc32(blk)
        crc = zng.crc32_combine(crc, c, len(blk))
        out.append(comp)
        prev = blk
        total += len(blk)
    assert zlib.decompressobj(-15).decompress(b"".join(out)) == fastq[:total]
    assert crc == zlib.crc32(fastq[:total])
    small = zng._ParallelCompress(1000, 6)
    with pytest.raises(OverflowError, match="Compressed output exceeds buffer size"):
        small.compress_and_crc(os.urandom(70_000), b"")
    with pytest.raises(zng.error, match="Bad compression level"):
        zng._ParallelCompress(1000, 42)


# ---------------------------------------------------------------- gzip_ng_threaded (tests/test_gzip_ng_threaded.py)
@pytest.mark.parametrize(["mode", "threads"], list(itertools.product(["wb", "wt"], [1, 3, -1])))
def test_threaded_write(gzt, fastq, tmp_path, mode, threads):
    target = tmp_path / "out.gz"
    text = fastq.decode("ascii")
    with gzt.open(target, mode, threads=threads, block_size=8 * 1024) as out_file:
        for i in range(0, len(fastq), 128 * 1024):
            out_file.write(fastq[i:i + 128 * 1024] if "b" in mode else text[i:i + 128 * 1024])
    with gzip.open(target, "rb") as f:
        assert f.read() == fastq
    raw = target.read_bytes()
    assert raw[:10] == b"\x1f\x8b\x08\x00\x00\x00\x00\x00\xff\x00"      # the reference's header bytes


def test_threaded_write_default_block_size_and_level1(gzt, tmp_path):
    data = synthdata.make("fastq", 9_000_000, 5).tobytes()
    target = tmp_path / "big.gz"
    with gzt.open(target, "wb", compresslevel=1, threads=4) as f:
        for i in range(0, len(data), 1 << 20):
            f.write(data[i:i + (1 << 20)])
    blob = target.read_bytes()
    assert gzip.decompress(blob) == data
    assert blob[9] == 4
    # 2-member file, the second one empty; the block index for the GPU reader follows as further
    # empty members (extra fields only), located by the fixed-size last member
    empty = b"\x1f\x8b\x08\x00\x00\x00\x00\x00\xff\x04\x03\x00" + b"\x00" * 8
    magic, version, start, raw_n, index_off, nblocks = gzt._LOCATOR.unpack(
        gzt._parse_extra_member(blob[-gzt._LOCATOR_MEMBER:], 0)[0])
    assert (magic, version, start, raw_n, nblocks) == (b"B2DX", 1, 0, len(data), (len(data) + 131071) // 131072)
    assert blob[:index_off].endswith(empty)


def test_threaded_write_without_index_ends_like_the_reference(gzt, tmp_path, monkeypatch):
    monkeypatch.setattr(gzt, "_WRITE_INDEX", False)
    data = synthdata.make("fastq", 1_000_000, 5).tobytes()
    target = tmp_path / "plain.gz"
    with gzt.open(target, "wb", compresslevel=6, threads=2) as f:
        f.write(data)
    blob = target.read_bytes()
    assert gzip.decompress(blob) == data
    assert blob.endswith(b"\x1f\x8b\x08\x00\x00\x00\x00\x00\xff\x00\x03\x00" + b"\x00" * 8)


def test_threaded_open_no_threads(gzt, tmp_path):
    p = tmp_path / "x.gz"
    p.write_bytes(gzip.compress(b"abc"))
    f = gzt.open(p, "rb", threads=0)
    assert isinstance(f, gzip.GzipFile)
    f.close()


def test_threaded_open_not_a_file_or_pathlike(gzt):
    with pytest.raises(TypeError, match="str") as e:
        gzt.open((1, 2, 3))
    e.match("bytes")
    e.match("file")


def test_threaded_read(gzt, fastq, tmp_path):
    p = tmp_path / "r.gz"
    p.write_bytes(gzip.compress(fastq[:500_000]))
    with gzt.open(p, "rb") as f:
        assert f.read() == fastq[:500_000]
        assert not f.writable()


@pytest.mark.timeout(20)
@pytest.mark.parametrize("threads", [1, 3])
def test_threaded_write_oversized_block_no_error(gzt, tmp_path, threads):
    data = os.urandom(1024 * 63)
    p = tmp_path / "o.gz"
    with gzt.open(p, "wb", compresslevel=3, threads=threads, block_size=8 * 1024) as w:
        w.write(data)
    assert gzip.decompress(p.read_bytes()) == data


@pytest.mark.timeout(20)
@pytest.mark.parametrize("threads", [1, 3])
def test_threaded_write_error(gzt, threads):
    f = gzt._ThreadedGzipWriter(io.BytesIO(), level=3, threads=threads, block_size=8 * 1024)
    f.input_queues[0].put((os.urandom(1024 * 64), b""))
    with pytest.raises(OverflowError, match="Compressed output exceeds buffer size"):
        f.close()


@pytest.mark.parametrize("threads", [1, 3])
def test_close_writer_twice_and_write_after_close(gzt, threads):
    f = gzt._ThreadedGzipWriter(io.BytesIO(), threads=threads)
    f.close()
    assert f.closed
    f.close()
    with pytest.raises(ValueError, match="closed"):
        f.write(b"abc")


def test_writer_argument_errors(gzt, zng, tmp_path):
    with pytest.raises(zng.error, match="Bad compression level"):
        gzt.open(tmp_path / "a.gz", mode="wb", compresslevel=42)
    with pytest.raises(ValueError, match="at least 1") as e:
        gzt._ThreadedGzipWriter(io.BytesIO(), threads=0)
    e.match("threads")
    with gzt.open(io.BytesIO(), "wb") as f:
        assert not f.readable()


def test_append_modes(gzt, tmp_path):
    p = tmp_path / "t.gz"
    with gzt.open(p, "wb") as f:
        f.write(b"AB")
    with gzt.open(p, mode="ab") as f:
        f.write(b"CD")
    assert gzip.decompress(p.read_bytes()) == b"ABCD"
    q = tmp_path / "t2.gz"
    with gzt.open(q, "wt") as f:
        f.write("AB")
    with gzt.open(q, mode="at") as f:
        f.write("CD")
    with gzip.open(q, "rt") as f:
        assert f.read() == "ABCD"


def test_streams_passed_in_are_not_closed(gzt):
    s = io.BytesIO()
    with gzt.open(s, "wb") as f:
        f.write(b"thisisatest")
    assert not s.closed
    assert gzip.decompress(s.getvalue()) == b"thisisatest"


@pytest.mark.parametrize("threads", [1, 2])
def test_flush(gzt, tmp_path, threads):
    p = tmp_path / "f.gz"
    with gzt.open(p, "wb", threads=threads) as f:
        f.write(b"1")
        f.flush()
        assert gzip.decompress(p.read_bytes()) == b"1"
        f.write(b"2")
        f.flush()
        assert gzip.decompress(p.read_bytes()) == b"12"
        f.write(b"3")
        f.flush()
        assert gzip.decompress(p.read_bytes()) == b"123"
    assert gzip.decompress(p.read_bytes()) == b"123"


@pytest.mark.timeout(60)
def test_program_can_exit_on_error(tmp_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    prog = tmp_path / "no_context_manager.py"
    out = tmp_path / "output.gz"
    prog.write_text(
        "import sys\n"
        f"sys.path.insert(0, {os.path.join(root, 'python-zlib-ng_b200')!r})\n"
        "from zlib_ng_b200 import gzip_ng_threaded\n"
        f"f = gzip_ng_threaded.open({str(out)!r}, mode='wb', threads=2)\n"
        "f.write(b'x' * 100000)\n"
        "raise Exception('Error')\n")
    r = subprocess.run([sys.executable, str(prog)], capture_output=True, timeout=50)
    assert r.returncode == 1 and b"Exception: Error" in r.stderr


@pytest.mark.parametrize("devices", ["0,0", "0,0,0", "0,1"])
def test_one_file_over_several_gpu_contexts(gzt, tmp_path, monkeypatch, devices):
    """The writer's multi-device path.  "0,0" puts two contexts on device 0, so a one-GPU box runs the same
    shard / concatenate / crc32_combine code as a box with several GPUs ("0,1" when there are two)."""
    from zlib_ng_b200 import _native
    if devices == "0,1" and _native.lib().b2d_device_count() < 2:
        pytest.skip("one GPU on this box")
    monkeypatch.setenv("B2D_DEVICES", devices)
    monkeypatch.setattr(gzt, "_STREAM_BATCH", 8 << 20)
    data = synthdata.make("mixed", 70_000_000, 9).tobytes()
    p = tmp_path / "m.gz"
    with gzt.open(p, "wb", threads=8, block_size=1 << 20) as f:
        assert len(f.raw._contexts[0].devices) == len(devices.split(","))
        for off in range(0, len(data), 9_000_001):
            f.write(data[off:off + 9_000_001])
    blob = p.read_bytes()
    assert gzip.decompress(blob) == data
    # the same bytes as one context on one device writes: sharding does not change the stream
    monkeypatch.setenv("B2D_DEVICES", "0")
    q = tmp_path / "one.gz"
    with gzt.open(q, "wb", threads=8, block_size=1 << 20) as f:
        f.write(data)
    assert q.read_bytes() == blob
    with gzt.open(p, "rb") as f:
        assert f.raw._gpu is not None and f.read() == data


@pytest.mark.parametrize("index", ["1", "0"])
def test_written_files_read_back_through_the_reference(gzt, fastq, tmp_path, monkeypatch, index):
    """north_star: "zlib_ng.decompress (the reference) must round-trip every output bit-exactly".  The reference's own
    reader (_GzipReader behind gzip_ng_threaded.open(..., "rb"), zlib_ngmodule.c:2426-2637, CRC / ISIZE check
    :2581-2599) and gzip_ng.decompress read what this writer wrote, with the block index and without
    (the reference's writer test decodes the same way: tests/test_gzip_ng_threaded.py:32-50)."""
    from helpers import reference_pkg
    ref = reference_pkg()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    ref_zlib, ref_gzip, ref_gzt = ref
    monkeypatch.setattr(gzt, "_WRITE_INDEX", index == "1")
    data = fastq * 3
    p = tmp_path / "r.gz"
    for level in (1, 6, 9):
        with gzt.open(p, "wb", compresslevel=level, threads=3, block_size=512 * 1024) as f:
            f.write(data)
        with ref_gzt.open(p, "rb") as f:
            assert f.read() == data
        assert ref_gzip.decompress(p.read_bytes()) == data
    with gzt.open(p, "wb", threads=2) as f:
        f.write(data[:1000])
        f.flush()
        f.write(data[1000:])
    with ref_gzt.open(p, "rb") as f:
        assert f.read() == data


def test_reference_decoder_on_every_api(zng, fastq):
    """The reference's zlib_ng.decompress / decompressobj as the decoder of compress(), compressobj() and
    compress_and_crc() output (all 21 wbits forms, zdict, flush modes)."""
    from helpers import reference
    ref = reference()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    for wbits in WBITS:
        for level in (1, 6):
            assert ref.decompress(zng.compress(fastq, level, wbits), wbits=wbits) == fastq
    zdict = fastq[:20_000]
    co = zng.compressobj(6, zng.DEFLATED, -15, zdict=zdict)
    blob = co.compress(fastq) + co.flush()
    assert ref.decompressobj(-15, zdict=zdict).decompress(blob) == fastq
    for mode in (zng.Z_SYNC_FLUSH, zng.Z_FULL_FLUSH, zng.Z_PARTIAL_FLUSH, zng.Z_BLOCK):
        co = zng.compressobj(6)
        blob = co.compress(fastq[:100_000]) + co.flush(mode) + co.compress(fastq[100_000:]) + co.flush()
        assert ref.decompress(blob) == fastq
    pc = zng._ParallelCompress(len(fastq) + len(fastq) // 10 + 500, 6)
    out, crc = pc.compress_and_crc(fastq, zdict)
    d = ref.decompressobj(-15, zdict=zdict)
    assert d.decompress(out + b"\x03\x00") == fastq and crc == ref.crc32(fastq)


def test_pinned_buffer_writes_are_not_copied_and_give_the_same_file(gzt, tmp_path, monkeypatch):
    monkeypatch.setattr(gzt, "_STREAM_BATCH", 8 << 20)
    data = synthdata.make("mixed", 40_000_000, 19).tobytes()
    a, b = tmp_path / "a.gz", tmp_path / "b.gz"
    with gzt.open(a, "wb", threads=4) as f:
        f.write(data)
    with gzt.pinned_buffer(len(data)) as pb:
        pb.view[:] = data
        with gzt.open(b, "wb", threads=4) as f:
            copies = []
            monkeypatch.setattr(gzt, "_parallel_copy", lambda *args: copies.append(args) or gzt._native.check(-2))
            assert f.write(pb.view) == len(data)
        assert not copies                                   # nothing went through the gather copy
    assert b.read_bytes() == a.read_bytes()
    assert gzip.decompress(b.read_bytes()) == data
    # a pinned write after an unaligned pageable one falls back to the gather path, and the other way round
    monkeypatch.undo()
    monkeypatch.setattr(gzt, "_STREAM_BATCH", 8 << 20)
    c = tmp_path / "c.gz"
    with gzt.pinned_buffer(20_000_000) as pb:
        pb.view[:] = data[5:20_000_005]
        with gzt.open(c, "wb", threads=4) as f:
            f.write(data[:5])
            f.write(pb.view)
            f.write(data[20_000_005:])
    assert gzip.decompress(c.read_bytes()) == data
    d = tmp_path / "d.gz"
    with gzt.pinned_buffer(16 << 20) as pb:
        pb.view[:] = data[:16 << 20]
        with gzt.open(d, "wb", threads=4) as f:
            f.write(pb.view)
            f.write(data[16 << 20:])
    assert d.read_bytes() == a.read_bytes()


# ---------------------------------------------------------------- gzip_ng
def test_gzip_ng_module(fastq, tmp_path):
    from zlib_ng_b200 import gzip_ng
    for level in (0, 1, 6, 9):
        blob = gzip_ng.compress(fastq[:300_000], level, mtime=1)
        assert gzip.decompress(blob) == fastq[:300_000]
        assert blob[4:8] == b"\x01\x00\x00\x00" and blob[9] == 255
    p = tmp_path / "g.gz"
    with gzip_ng.open(p, "wb", compresslevel=6) as f:
        for i in range(0, 2_000_000, 333_333):
            f.write(fastq[i:min(i + 333_333, 2_000_000)])
    with gzip.open(p, "rb") as f:
        assert f.read() == fastq[:2_000_000]
    with gzip_ng.open(p, "rb") as f:
        assert f.read(10) == fastq[:10]
    assert gzip_ng.decompress(p.read_bytes()) == fastq[:2_000_000]


def test_compressobj_copy_of_a_long_stream(zng, monkeypatch):
    """copy() while the stream gathers its input in page-locked memory (long streams switch to it): both objects go on
    independently and produce what one compress() call produces"""
    monkeypatch.setattr(zng, "_STREAM_THRESHOLD", 1 << 20)
    monkeypatch.setattr(zng, "_PIN_BYTES", 1 << 20)
    data = synthdata.make("mixed", 6_000_000, 23).tobytes()
    co = zng.compressobj(6, zng.DEFLATED, 31)
    head = b"".join(co.compress(data[i:i + 300_000]) for i in range(0, 3_300_000, 300_000))
    assert co._pin is not None and co._pin_fill > 0          # part of the input sits in the gather buffer
    twin = co.copy()
    a = head + co.compress(data[3_300_000:]) + co.flush()
    b = head + twin.compress(data[3_300_000:5_000_000]) + twin.compress(data[5_000_000:]) + twin.flush()
    assert a == b == zng.compress(data, 6, 31)
    assert zlib.decompress(a, 31) == data
    # flush modes in the middle of a long stream
    co = zng.compressobj(1)
    parts = [co.compress(data[:2_500_000]), co.flush(zng.Z_SYNC_FLUSH), co.compress(data[2_500_000:]), co.flush(zng.Z_FULL_FLUSH), co.flush()]
    assert zlib.decompress(b"".join(parts)) == data
