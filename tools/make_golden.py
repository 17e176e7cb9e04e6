#!/usr/bin/env python3
"""Generate tests/golden/*.json from the reference (run HERE, where
/root/reference exists; the GPU box only ever reads the committed JSON).

Two sources, kept apart in the output:
  * "kat"   rows lifted from the reference's own test tables
            (zlib-ng test/test_crc32.cc:27-181, test/test_adler32.cc:201-344,
            tests/test_zlib_compliance.py:94-127)
  * "ref"   values produced by running the reference itself (oracle/_ref,
            built by oracle/Makefile) on inputs that the tests can regenerate
            from synthdata.py seeds.
"""
import base64
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
ZNG_TEST = REF + "/src/zlib_ng/zlib-ng/test"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "pkg"))

import numpy as np  # noqa: E402
import synthdata  # noqa: E402
from zlib_ng import zlib_ng  # noqa: E402


def _c_unescape(s: str) -> bytes:
    return s.encode("latin1").decode("unicode_escape").encode("latin1")


def parse_rows(path, long_string=None):
    text = open(path, encoding="latin1").read()
    body = text[text.index("tests[] = {"):]
    body = body[:body.index("};")]
    rows = []
    # {start, (const uint8_t *)<buf>, len, expect}
    pat = re.compile(r"\{\s*(0x[0-9a-fA-F]+)\s*,\s*\(const uint8_t \*\)\s*((?:\"(?:[^\"\\]|\\.)*\"\s*)+|0x0|long_string)\s*,"
                     r"\s*(\d+)\s*,\s*(0x[0-9a-fA-F]+)\s*\}", re.S)
    for m in pat.finditer(body):
        start, buf, ln, expect = m.groups()
        ln = int(ln)
        if buf == "0x0":
            data = None
        elif buf == "long_string":
            data = long_string
        else:
            data = b"".join(_c_unescape(x) for x in re.findall(r"\"((?:[^\"\\]|\\.)*)\"", buf))
            # C string literals carry an implicit NUL the tables rely on for len == strlen+1
            data = (data + b"\0")[:ln] if ln > len(data) else data[:ln]
        rows.append({"start": int(start, 16), "null": data is None, "len": ln,
                     "data": None if data is None else base64.b64encode(data).decode(),
                     "expect": int(expect, 16)})
    return rows


def parse_long_string(path):
    text = open(path, encoding="latin1").read()
    body = text[text.index("long_string[5552] = {"):]
    body = body[body.index("{") + 1:body.index("};")]
    chars = re.findall(r"'(\\.|[^'\\])'", body)
    out = bytes(_c_unescape(c)[0] for c in chars)
    assert len(out) == 5552, len(out)
    return out


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    long_string = parse_long_string(ZNG_TEST + "/test_adler32.cc")
    crc_rows = parse_rows(ZNG_TEST + "/test_crc32.cc")
    adl_rows = parse_rows(ZNG_TEST + "/test_adler32.cc", long_string)
    assert len(crc_rows) > 100 and len(adl_rows) > 100, (len(crc_rows), len(adl_rows))
    # sanity: the reference must agree with its own tables
    for r in crc_rows:
        if not r["null"] and r["len"]:
            assert zlib_ng.crc32(base64.b64decode(r["data"]), r["start"]) == r["expect"]
    for r in adl_rows:
        if not r["null"]:
            assert zlib_ng.adler32(base64.b64decode(r["data"]), r["start"]) == r["expect"], r
    py_kat = {
        # tests/test_zlib_compliance.py:94-109
        "crc32": [["penguin", 0, 0x0e5c1a120], ["penguin", 1, 0x43b6aa94],
                  ["abcdefghijklmnop", 0, 2486878355], ["spam", 0, 1138425661]],
        "adler32": [["penguin", 0, 0x0bcf02f6], ["penguin", 1, 0x0bd602f7],
                    ["spam", 1, 72286642]],
    }
    json.dump({"source": "zlib-ng test/test_crc32.cc, test/test_adler32.cc; tests/test_zlib_compliance.py",
               "crc32": crc_rows, "adler32": adl_rows, "python": py_kat},
              open(os.path.join(out_dir, "checksum_kat.json"), "w"), indent=0)

    # ---- values produced by the reference itself on regenerable inputs ----
    ref = {"checksums": [], "combine": [], "sizes": [], "streams": []}
    rng = np.random.Generator(np.random.PCG64(7))
    for kind, seed, n in (("fastq", 1, 1 << 20), ("mixed", 2, 3 << 20), ("mixed", 5, 70001),
                          ("fastq", 9, 5553), ("mixed", 11, 1)):
        data = synthdata.make(kind, n, seed).tobytes()
        for start in (0, 1, 0xDEADBEEF, 0xFFFFFFFF):
            ref["checksums"].append({"kind": kind, "seed": seed, "n": n, "start": start,
                                     "crc32": zlib_ng.crc32(data, start),
                                     "adler32": zlib_ng.adler32(data, start)})
    for _ in range(64):
        c1, c2 = (int(x) for x in rng.integers(0, 1 << 32, 2))
        ln = int(rng.choice([0, 1, 2, 5552, 65536, 131072, (1 << 31) + 5, (1 << 33) + 12345,
                             int(rng.integers(0, 1 << 40))]))
        ref["combine"].append({"crc1": c1, "crc2": c2, "len2": ln,
                               "expect": zlib_ng.crc32_combine(c1, c2, ln)})

    def ref_blocks(data, level, bs):
        c = zlib_ng._ParallelCompress(bs + max(bs // 10, 500), level)
        total, prev = 0, b""
        for i in range(0, len(data), bs):
            blk = data[i:i + bs]
            out, _ = c.compress_and_crc(blk, prev[-32768:])
            total += len(out)
            prev = blk
        return total

    for kind, seed, n in (("fastq", 1, 4 << 20), ("mixed", 2, 8 << 20), ("fastq", 3, 1 << 20),
                          ("mixed", 4, 1 << 20)):
        data = synthdata.make(kind, n, seed).tobytes()
        for level in (1, 2, 3, 4, 5, 6, 7, 8, 9):
            ref["sizes"].append({"kind": kind, "seed": seed, "n": n, "level": level,
                                 "block_size": 131072,
                                 "threaded_blocks": ref_blocks(data, level, 131072),
                                 "single_stream": len(zlib_ng.compress(data, level, -15))})
    # small reference-made streams: pin the oracle's inflate against real zlib-ng output
    for kind, seed, n in (("fastq", 21, 40000), ("mixed", 22, 150000), ("mixed", 23, 9000)):
        data = synthdata.make(kind, n, seed).tobytes()
        for level in (0, 1, 6, 9):
            comp = zlib_ng.compress(data, level, -15)
            ref["streams"].append({"kind": kind, "seed": seed, "n": n, "level": level, "wbits": -15,
                                   "crc32": zlib_ng.crc32(data),
                                   "stream": base64.b64encode(comp).decode()})
        c = zlib_ng._ParallelCompress(n + 1000, 6)
        comp, crc = c.compress_and_crc(data[n // 2:], data[:n // 2][-32768:])
        ref["streams"].append({"kind": kind, "seed": seed, "n": n, "level": 6, "wbits": -15,
                               "dict_split": n // 2, "crc32": crc,
                               "stream": base64.b64encode(comp).decode()})
    json.dump(ref, open(os.path.join(out_dir, "reference_vectors.json"), "w"), indent=0)
    print("crc rows", len(crc_rows), "adler rows", len(adl_rows), "ref checksums", len(ref["checksums"]),
          "combine", len(ref["combine"]), "sizes", len(ref["sizes"]), "streams", len(ref["streams"]))


if __name__ == "__main__":
    main()
