#!/usr/bin/env python3
"""tests/golden/fullsize_ratios.json: the REFERENCE's compressed size of the 256 MiB base segments the full-size GPU tests
tile to BASELINE.json's sizes (tests/test_gpu_fullsize.py), produced by running the reference itself (oracle/_ref) the way
its threaded writer does — blocks of 1 MiB through _ParallelCompress.compress_and_crc, each primed with the 32 KiB before
it (src/zlib_ng/gzip_ng_threaded.py:360-380).  Run HERE (the GPU box only reads the committed JSON)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "pkg"))
import synthdata  # noqa: E402
from zlib_ng import zlib_ng  # noqa: E402

MIB = 1 << 20
rows = []
for kind, seed, level in (("fastq", 1, 1), ("mixed", 2, 6)):
    n, bs = 256 * MIB, MIB
    data = synthdata.make(kind, n, seed).tobytes()
    c = zlib_ng._ParallelCompress(bs + max(bs // 10, 500), level)
    total, prev = 0, b""
    for i in range(0, n, bs):
        blk = data[i:i + bs]
        out, _ = c.compress_and_crc(blk, prev[-32768:])
        total += len(out)
        prev = blk
    rows.append({"kind": kind, "seed": seed, "base_bytes": n, "level": level, "block_size": bs, "reference_bytes": total,
                 "reference_ratio": total / n})
    print(rows[-1])
json.dump({"source": "oracle/_ref (the reference's _ParallelCompress), tools/make_golden_fullsize.py", "rows": rows},
          open(os.path.join(ROOT, "tests", "golden", "fullsize_ratios.json"), "w"), indent=1)
