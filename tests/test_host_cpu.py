"""CPU-only tests: the C-ABI library loads and exports what include/b2d.h
declares, its pure-host functions agree with the oracle and the golden
vectors, and the CPU twin of the kernels (the byte-for-byte model the GPU tests
compare against) round-trips and stays within the ratio tolerance of the
reference's own sizes (tests/golden/reference_vectors.json, produced by the
reference built from /root/reference)."""
import os
import re
import zlib

import numpy as np
import pytest

import synthdata
from helpers import ROOT, golden, oracle, orc_inflate, twin_deflate


def _native():
    from zlib_ng_b200 import _native
    return _native


def test_library_exports_every_declared_symbol():
    nat = _native()
    hdr = open(os.path.join(ROOT, "include", "b2d.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b2d_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    assert declared == set(nat.SIGNATURES), declared ^ set(nat.SIGNATURES)
    L = nat.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_no_cpu_fallback_without_a_device():
    nat = _native()
    if nat.lib().b2d_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(nat.NativeError) as e:
        nat.Context(0)
    assert "no CPU fallback" in str(e.value)
    from zlib_ng_b200 import zlib_ng
    with pytest.raises(zlib_ng.error):
        zlib_ng.crc32(b"abc")
    with pytest.raises(zlib_ng.error):
        zlib_ng.compress(b"abc")


def test_host_combine_functions():
    L = _native().lib()
    O = oracle()
    for r in golden("reference_vectors.json")["combine"]:
        assert L.b2d_crc32_combine(r["crc1"], r["crc2"], r["len2"]) == r["expect"]
    rng = np.random.default_rng(2)
    data = rng.integers(0, 256, 200_000, dtype=np.uint8).tobytes()
    for cut in (0, 1, 5552, 65521, 100_000, 199_999, 200_000):
        a, b = data[:cut], data[cut:]
        assert L.b2d_crc32_combine(zlib.crc32(a), zlib.crc32(b), len(b)) == zlib.crc32(data)
        assert L.b2d_adler32_combine(zlib.adler32(a), zlib.adler32(b), len(b)) == zlib.adler32(data)
        assert L.b2d_adler32_combine(zlib.adler32(a), zlib.adler32(b), len(b)) == \
            O.orc_adler32_combine(zlib.adler32(a), zlib.adler32(b), len(b))
    assert L.b2d_adler32_combine(1, 1, -1) == 0xFFFFFFFF


def test_deflate_bound_covers_stored_worst_case():
    L = _native().lib()
    for n in (0, 1, 32767, 32768, 32769, 131072, 131073, 1 << 20, (1 << 30) + 7):
        nsub, nblk = -(-n // 32768), max(-(-n // 131072), 1)
        assert L.b2d_deflate_bound(n) >= n + 5 * nsub + 5 * nblk + 2


def test_python_layer_argument_handling_without_gpu():
    from zlib_ng_b200 import zlib_ng
    assert zlib_ng.crc32_combine(0, 0, 0) == 0
    with pytest.raises(TypeError):
        zlib_ng.crc32_combine("a", 1, 1)
    assert zlib_ng._wrapper(15) == ("zlib", 15) and zlib_ng._wrapper(-9) == ("raw", 9)
    assert zlib_ng._wrapper(31) == ("gzip", 15) and zlib_ng._wrapper(8) == ("zlib", 9)
    with pytest.raises(ValueError):
        zlib_ng._wrapper(0)
    with pytest.raises(zlib_ng.error):
        zlib_ng._level(10)
    assert zlib_ng._level(-1) == 6
    assert zlib_ng._gzip_header(9)[8:] == b"\x02\x03" and zlib_ng._gzip_header(1)[8] == 4
    for lvl, flg in ((1, 0x01), (5, 0x5E), (6, 0x9C), (9, 0xDA)):
        assert zlib_ng._zlib_header(lvl, 15, b"") == bytes([0x78, flg])
    # inflate half is served by CPython's zlib
    assert zlib_ng.decompress(zlib.compress(b"abc")) == b"abc"


# ---------------------------------------------------------------- the twin (model of the kernels)
@pytest.mark.parametrize("kind,n,seed", [("fastq", 300_000, 1), ("mixed", 1_000_000, 2), ("mixed", 131_073, 3)])
@pytest.mark.parametrize("level", [0, 1, 3, 6, 9])
def test_twin_round_trips(kind, n, seed, level):
    data = synthdata.make(kind, n, seed).tobytes()
    out = twin_deflate(data, level)
    assert out.endswith(b"\x00\x00\xff\xff")
    assert orc_inflate(out, n, need_final=False) == data
    assert zlib.decompressobj(-15).decompress(out) == data


def test_twin_with_dictionary():
    data = synthdata.make("fastq", 400_000, 5).tobytes()
    zdict, body = data[:150_000], data[150_000:]
    out = twin_deflate(body, 6, zdict)
    assert zlib.decompressobj(-15, zdict=zdict[-32768:]).decompress(out) == body
    assert orc_inflate(out, len(body), zdict=zdict, need_final=False) == body
    assert len(out) < len(twin_deflate(body, 6))


def test_twin_degenerate_inputs():
    for data in (b"\0" * 200_000, b"a", b"ab" * 40_000, os.urandom(100_000), bytes(range(256)) * 300):
        for level in (1, 6):
            assert zlib.decompressobj(-15).decompress(twin_deflate(data, level)) == data
    rnd = os.urandom(300_000)
    assert len(twin_deflate(rnd, 6)) <= len(rnd) + 5 * 10 + 5 * 3      # stored fallback


@pytest.mark.parametrize("row", [r for r in golden("reference_vectors.json")["sizes"] if r["n"] <= (4 << 20)],
                         ids=lambda r: f"{r['kind']}{r['seed']}-L{r['level']}")
def test_ratio_vs_reference_golden(row):
    """size(ours, level L) <= 1.03 x size(reference, level L, same 128 KiB blocks) — BASELINE.json's tolerance."""
    data = synthdata.make(row["kind"], row["n"], row["seed"]).tobytes()
    ours = len(twin_deflate(data, row["level"]))
    assert ours <= 1.03 * row["threaded_blocks"], (ours, row["threaded_blocks"], ours / row["threaded_blocks"])
    assert ours <= 1.03 * row["single_stream"]


@pytest.mark.parametrize("level", [1, 6])
def test_twin_strategies(level):
    """compressobj(strategy=...) semantics of the reference (deflate_huff.c, deflate_rle.c, trees.c:677, the
    Z_FILTERED rule of deflate_slow.c) in the CPU twin the GPU is compared with."""
    text = synthdata.make("mixed", 300_000, 3).tobytes()
    runs = b"".join(bytes([i % 251]) * 700 for i in range(300))
    periodic = b"abcdefgh" * 30_000
    size = {}
    for name, data in (("text", text), ("runs", runs), ("periodic", periodic)):
        for strat in range(5):
            out = twin_deflate(data, level | (strat << 8))
            assert zlib.decompressobj(-15).decompress(out) == data
            size[name, strat] = len(out)
            if strat == 4:        # Z_FIXED: first block is stored (00) or fixed (01), never dynamic (10)
                assert (out[0] >> 1) & 3 in (0, 1)
    assert size["runs", 2] > 20 * size["runs", 0]            # Huffman only: runs are not found
    assert size["runs", 3] <= 1.1 * size["runs", 0]          # RLE finds them
    assert size["periodic", 3] > 20 * size["periodic", 0]    # ... but not a period of 8
    assert size["text", 1] != size["text", 0]                # filtered drops matches shorter than 6
    assert size["text", 4] >= size["text", 0]


def test_index_members_are_ordinary_empty_gzip_members():
    """The block index the threaded writer appends (gzip_ng_threaded._extra_member): empty members whose header has
    an extra field, so every gzip reader skips them; the fixed-size locator parses back."""
    import gzip
    import struct
    from zlib_ng_b200 import gzip_ng_threaded as gzt
    sizes = list(range(1000, 1300))
    idx = gzt._extra_member(struct.pack("<BBHI", 1, 0, 0, len(sizes)) + struct.pack(f"<{len(sizes)}I", *sizes))
    loc = gzt._extra_member(gzt._LOCATOR.pack(gzt._LOCATOR_MAGIC, 1, 0, 12345, 777, len(sizes)))
    assert len(loc) == gzt._LOCATOR_MEMBER
    blob = gzip.compress(b"payload") + idx + loc
    assert gzip.decompress(blob) == b"payload"
    payload, nxt = gzt._parse_extra_member(idx + loc, 0)
    assert struct.unpack_from(f"<{len(sizes)}I", payload, 8) == tuple(sizes) and nxt == len(idx)
    assert gzt._LOCATOR.unpack(gzt._parse_extra_member(loc, 0)[0])[3] == 12345
    assert gzt._parse_extra_member(gzip.compress(b"x"), 0) is None


def test_host_copy_needs_no_device():
    """b2d_host_copy is plain host code (the writer's copy into its pinned batch): any thread count, any size."""
    import ctypes
    import numpy as np
    from zlib_ng_b200 import _native
    L = _native.lib()
    rng = np.random.default_rng(1)
    for n, threads in ((0, 4), (1, 1), (4097, 8), (9 << 20, 3), (33 << 20, 8), (17 << 20, 64)):
        src = rng.integers(0, 256, n, dtype=np.uint8)
        dst = np.zeros(n + 8, dtype=np.uint8)
        assert L.b2d_host_copy(dst.ctypes.data_as(ctypes.c_void_p), src.ctypes.data_as(ctypes.c_void_p), n, threads) == 0
        assert (dst[:n] == src).all() and not dst[n:].any()


def test_reader_without_a_device_reads_indexed_files_through_cpython(tmp_path):
    """An indexed file on a machine without a GPU: the reader finds the index, cannot open a device and reads the
    file the ordinary way (reading was never the accelerated path: the reference reads on one thread too)."""
    import gzip
    import struct
    from zlib_ng_b200 import _native, gzip_ng_threaded as gzt
    if _native.lib().b2d_device_count() > 0:
        pytest.skip("a CUDA device is present: the GPU tests cover the indexed path")
    data = bytes(range(256)) * 300
    member = gzip.compress(data, mtime=0)
    assert member[:4] == b"\x1f\x8b\x08\x00"
    empty = struct.pack("<BBBBIBB", 0x1F, 0x8B, 8, 0, 0, 0xFF, 0) + b"\x03\x00" + b"\x00" * 8
    sizes = [len(member) - 10 - 8 - 2]
    index_off = len(member) + len(empty)
    blob = (member + empty
            + gzt._extra_member(struct.pack("<BBHI", 1, 0, 0, 1) + struct.pack("<I", *sizes))
            + gzt._extra_member(gzt._LOCATOR.pack(gzt._LOCATOR_MAGIC, 1, 0, len(data), index_off, 1)))
    p = tmp_path / "indexed.gz"
    p.write_bytes(blob)
    assert gzip.decompress(blob) == data
    with gzt.open(p, "rb") as f:
        assert f.raw._gpu is None
        assert f.read() == data


def _bgzf(data, blk=65280):
    import struct
    out = []
    for piece in [data[o:o + blk] for o in range(0, len(data), blk)] + [b""]:
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = c.compress(piece) + c.flush()
        out.append(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", len(body) + 25) + body
                   + struct.pack("<II", zlib.crc32(piece), len(piece)))
    return b"".join(out)


def test_bgzf_scan_is_plain_host_code(tmp_path):
    """b2d_bgzf_scan finds the members of a BGZF file through their BC subfields (the layout of the reference's fixture
    tests/data/test.fastq.bgzip.gz); without a device the reader hands BGZF files to CPython like any other gzip file."""
    import ctypes
    import gzip
    import synthdata
    from zlib_ng_b200 import _native, gzip_ng_threaded
    L = _native.lib()
    data = synthdata.make("fastq", 700_000, 3).tobytes()
    blob = _bgzf(data)
    n = len(blob)
    m = n // 28 + 16
    po, pl = (ctypes.c_uint64 * m)(), (ctypes.c_uint32 * m)()
    rl, cr = (ctypes.c_uint32 * m)(), (ctypes.c_uint32 * m)()
    cnt, used = ctypes.c_uint64(0), ctypes.c_uint64(0)
    assert L.b2d_bgzf_scan(blob, n, m, po, pl, rl, cr, ctypes.byref(cnt), ctypes.byref(used)) == 0
    assert used.value == n and cnt.value == -(-len(data) // 65280) + 1 and sum(rl[:cnt.value]) == len(data)
    off = 0
    for i in range(cnt.value):
        piece = zlib.decompressobj(-15).decompress(blob[po[i]:po[i] + pl[i]])
        assert piece == data[off:off + rl[i]] and zlib.crc32(piece) == cr[i]
        off += rl[i]
    # a buffer cut inside a member stops at the last complete one; data that is not BGZF yields no members
    assert L.b2d_bgzf_scan(blob, n - 40, m, po, pl, rl, cr, ctypes.byref(cnt), ctypes.byref(used)) == 0 and used.value < n - 40
    plain = gzip.compress(data[:1000])
    assert L.b2d_bgzf_scan(plain, len(plain), m, po, pl, rl, cr, ctypes.byref(cnt), ctypes.byref(used)) == 0 and cnt.value == 0
    if L.b2d_device_count() == 0:
        p = tmp_path / "x.bgzf.gz"
        p.write_bytes(blob)
        with gzip_ng_threaded.open(p, "rb") as f:
            assert f.raw._gpu is None and f.read() == data


def test_reader_windows_cover_every_block_once_and_ramp_up():
    """the reader's GPU calls: contiguous, nothing left out, at most `window` blocks each, the first ones smaller"""
    from zlib_ng_b200.gzip_ng_threaded import _reader_windows
    for nblocks, window in ((0, 4096), (1, 4096), (1023, 4096), (1024, 4096), (4096, 4096), (32768, 4096), (32769, 4096), (7, 1), (100, 3)):
        wins = _reader_windows(nblocks, window)
        assert sum(nb for _, nb in wins) == nblocks
        pos = 0
        for first, nb in wins:
            assert first == pos and 1 <= nb <= window
            pos += nb
        if nblocks >= 4 * window and window >= 4:
            assert [nb for _, nb in wins[:3]] == [window // 4, window // 2, window]
