"""Pin the oracle (oracle/oracle.c) against the reference's own vectors and,
where oracle/_ref is present, against the reference itself.  CPU only."""
import zlib

import numpy as np
import pytest

import synthdata
from helpers import b64, golden, oracle, orc_inflate, reference


def test_crc32_kat_rows():
    L = oracle()
    rows = golden("checksum_kat.json")["crc32"]
    assert len(rows) >= 140
    for r in rows:
        if r["null"]:
            continue
        data = b64(r["data"])
        got = L.orc_crc32(r["start"], data, len(data)) if r["len"] else r["start"]
        assert got == r["expect"], r


def test_adler32_kat_rows():
    L = oracle()
    rows = golden("checksum_kat.json")["adler32"]
    assert len(rows) >= 140
    for r in rows:
        if r["null"]:
            continue
        data = b64(r["data"])
        assert L.orc_adler32(r["start"], data, len(data)) == r["expect"], r


def test_python_level_kats():
    L = oracle()
    k = golden("checksum_kat.json")["python"]
    for s, start, expect in k["crc32"]:
        assert L.orc_crc32(start, s.encode(), len(s)) == expect
    for s, start, expect in k["adler32"]:
        assert L.orc_adler32(start, s.encode(), len(s)) == expect


def test_reference_generated_checksums():
    L = oracle()
    for r in golden("reference_vectors.json")["checksums"]:
        data = synthdata.make(r["kind"], r["n"], r["seed"]).tobytes()
        assert L.orc_crc32(r["start"], data, len(data)) == r["crc32"]
        assert L.orc_adler32(r["start"], data, len(data)) == r["adler32"]


def test_crc32_combine_vectors():
    L = oracle()
    for r in golden("reference_vectors.json")["combine"]:
        assert L.orc_crc32_combine(r["crc1"], r["crc2"], r["len2"]) == r["expect"], r


def test_combine_identities():
    # tests/test_zlib_compliance.py:111-120 in the reference
    L = oracle()
    rng = np.random.default_rng(3)
    data = rng.integers(0, 256, 100_000, dtype=np.uint8).tobytes()
    for cut in (0, 1, 5551, 5552, 5553, 65521, 99_999, 100_000):
        a, b = data[:cut], data[cut:]
        whole = L.orc_crc32(0, data, len(data))
        assert L.orc_crc32_combine(L.orc_crc32(0, a, len(a)), L.orc_crc32(0, b, len(b)), len(b)) == whole
        assert whole == zlib.crc32(data)
        wa = L.orc_adler32(1, data, len(data))
        assert L.orc_adler32_combine(L.orc_adler32(1, a, len(a)), L.orc_adler32(1, b, len(b)), len(b)) == wa
        assert wa == zlib.adler32(data)


def test_inflate_on_reference_streams():
    for r in golden("reference_vectors.json")["streams"]:
        data = synthdata.make(r["kind"], r["n"], r["seed"]).tobytes()
        stream = b64(r["stream"])
        if "dict_split" in r:
            cut = r["dict_split"]
            out = orc_inflate(stream, len(data), zdict=data[:cut][-32768:], need_final=False)
            assert out == data[cut:]
            assert oracle().orc_crc32(0, out, len(out)) == r["crc32"]
        else:
            out = orc_inflate(stream, len(data))
            assert out == data
            assert zlib.decompress(stream, -15) == data


def test_inflate_rejects_garbage():
    with pytest.raises(ValueError):
        orc_inflate(b"\x07\xff\xff\xff", 100)       # BTYPE=3
    with pytest.raises(ValueError):
        orc_inflate(b"\x01\x05\x00\x00\x00abcde", 100)  # bad NLEN


@pytest.mark.skipif(reference() is None, reason="oracle/_ref not built")
def test_oracle_vs_live_reference():
    ref = reference()
    L = oracle()
    rng = np.random.default_rng(11)
    for n in (0, 1, 7, 8, 63, 64, 65, 4095, 5552, 5553, 65536, 1 << 20):
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        for start in (0, 1, 0xFFFFFFFF, 123456789):
            assert L.orc_crc32(start, data, n) == ref.crc32(data, start)
            assert L.orc_adler32(start, data, n) == ref.adler32(data, start)
    for kind, seed in (("fastq", 31), ("mixed", 32)):
        data = synthdata.make(kind, 300_000, seed).tobytes()
        for level in range(0, 10):
            assert orc_inflate(ref.compress(data, level, -15), len(data)) == data


def test_fullsize_ratio_fixture_is_complete():
    """the full-size GPU tests take the reference's sizes from this fixture (tools/make_golden_fullsize.py)"""
    import json
    import os
    rows = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_ratios.json")))["rows"]
    got = {(r["kind"], r["seed"], r["level"]): r for r in rows}
    assert set(got) == {("fastq", 1, 1), ("mixed", 2, 6)}
    for r in rows:
        assert r["base_bytes"] == 256 << 20 and r["block_size"] == 1 << 20
        assert abs(r["reference_ratio"] - r["reference_bytes"] / r["base_bytes"]) < 1e-12
        assert 0.3 < r["reference_ratio"] < 0.8
