#!/usr/bin/env python3
"""Dev tool: end-to-end throughput of gzip_ng_threaded.open(..., "wb") (BASELINE.json configs[4] shape, scaled):
the user-facing call, host bytes in -> gzip bytes to a null sink, this repo vs the reference on the host cores.

usage: stream_bench.py [--gib 4] [--level 6] [--threads 16] [--block-mib 1] [--write-mib 64] [--ref] [--verify]
"""
import argparse
import io
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))

import numpy as np  # noqa: E402
import synthdata  # noqa: E402


class NullSink(io.RawIOBase):
    def __init__(self):
        self.n = 0

    def writable(self):
        return True

    def write(self, b):
        self.n += len(b)
        return len(b)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=4.0)
    ap.add_argument("--level", type=int, default=6)
    ap.add_argument("--threads", type=int, default=16)
    ap.add_argument("--block-mib", type=float, default=1.0)
    ap.add_argument("--write-mib", type=int, default=64)
    ap.add_argument("--kind", default="mixed")
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--verify", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    n = int(a.gib * (1 << 30))
    base = synthdata.make(a.kind, min(n, 256 << 20), 2)
    data = np.tile(base, (n + base.size - 1) // base.size)[:n]
    mv = memoryview(data)
    if a.ref:
        sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "pkg"))
        from zlib_ng import gzip_ng_threaded as gt
    else:
        from zlib_ng_b200 import gzip_ng_threaded as gt
    step = a.write_mib << 20
    for r in range(a.reps):
        sink = io.BytesIO() if a.verify else NullSink()
        t0 = time.perf_counter()
        with gt.open(sink, "wb", compresslevel=a.level, threads=a.threads, block_size=int(a.block_mib * (1 << 20))) as f:
            for off in range(0, n, step):
                f.write(mv[off:off + step])
        dt = time.perf_counter() - t0
        out_n = len(sink.getvalue()) if a.verify else sink.n
        print(f"rep{r}: {n / dt / 1e9:.2f} GB/s  ({dt * 1e3:.0f} ms)  ratio {out_n / n:.4f}", flush=True)
        if a.verify:
            import gzip
            ok = gzip.decompress(sink.getvalue()) == data.tobytes()
            print("round trip", ok)
            a.verify = False


if __name__ == "__main__":
    main()
