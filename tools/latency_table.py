#!/usr/bin/env python3
"""Small-input crossover (VERDICT r1 missing #5): microseconds per call of crc32 / adler32 / compress / compress_and_crc at
1 KiB ... 16 MiB, this library beside the reference (oracle/_ref) on the same box — the sizes of the reference's own
benchmark script (benchmark_scripts/benchmark.py:56-75) and above.  Prints one JSON object; DESIGN.md holds the table.
    python tools/latency_table.py > profiles/latency_r2.json"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "pkg"))
import synthdata  # noqa: E402
from zlib_ng_b200 import zlib_ng as ours  # noqa: E402

try:
    from zlib_ng import zlib_ng as ref  # type: ignore
except Exception:
    ref = None

SIZES = [1 << 10, 8 << 10, 64 << 10, 1 << 20, 16 << 20]
base = synthdata.make("fastq", 16 << 20, 5).tobytes()


def per_call_us(fn, budget=0.4):
    fn()                                   # warm
    n, t0 = 0, time.perf_counter()
    while True:
        fn()
        n += 1
        dt = time.perf_counter() - t0
        if dt > budget or n >= 2000:
            return dt / n * 1e6


res = {"sizes": SIZES, "rows": []}
for size in SIZES:
    data = base[:size]
    row = {"bytes": size}
    for name, mod in (("ours", ours), ("reference", ref)):
        if mod is None:
            continue
        pc = mod._ParallelCompress(size + size // 10 + 500, 6)
        row[name] = {
            "crc32_us": per_call_us(lambda: mod.crc32(data)),
            "adler32_us": per_call_us(lambda: mod.adler32(data)),
            "compress_L6_us": per_call_us(lambda: mod.compress(data, 6)),
            "compress_L1_us": per_call_us(lambda: mod.compress(data, 1)),
            "compress_and_crc_L6_us": per_call_us(lambda: pc.compress_and_crc(data, b"")),
        }
    res["rows"].append(row)
print(json.dumps(res))
