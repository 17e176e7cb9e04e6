#!/usr/bin/env python3
"""Dev tool: throughput of the one-shot / streaming Python calls on pageable host bytes."""
import os, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np
import synthdata
from zlib_ng_b200 import zlib_ng
n = int(float(sys.argv[1]) * (1 << 30)) if len(sys.argv) > 1 else 1 << 30
base = synthdata.make("mixed", min(n, 256 << 20), 2)
data = np.tile(base, (n + base.size - 1) // base.size)[:n].tobytes()
for name, fn in (("compress(wbits=31)", lambda: zlib_ng.compress(data, 6, 31)),
                 ("crc32", lambda: zlib_ng.crc32(data)), ("adler32", lambda: zlib_ng.adler32(data))):
    for r in range(3):
        t = time.perf_counter(); out = fn(); dt = time.perf_counter() - t
        print(f"{name} rep{r}: {n / dt / 1e9:.2f} GB/s", flush=True)
    if name.startswith("compress"):
        print("  ok", zlib.decompress(out, 31) == data, "ratio", len(out) / n)
co = zlib_ng.compressobj(6, zlib_ng.DEFLATED, 31)
t = time.perf_counter(); tot = 0
for off in range(0, n, 8 << 20):
    tot += len(co.compress(data[off:off + (8 << 20)]))
tot += len(co.flush()); dt = time.perf_counter() - t
print(f"compressobj 8 MiB writes: {n / dt / 1e9:.2f} GB/s ratio {tot / n:.4f}")
