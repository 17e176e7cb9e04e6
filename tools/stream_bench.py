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
    ap.add_argument("--read", action="store_true", help="write the file once (this repo's writer), then time open(..., 'rb') reads")
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
    if a.read:
        import tempfile
        from zlib_ng_b200 import gzip_ng_threaded as ours
        d = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
        path = os.path.join(d, "b2d_stream_bench.gz")
        with ours.open(path, "wb", compresslevel=a.level, threads=a.threads) as f:
            for off in range(0, n, step):
                f.write(mv[off:off + step])
        print(f"file {os.path.getsize(path) / n:.4f} of the input; reader = {'reference' if a.ref else 'this repo'}", flush=True)
        import zlib
        want = zlib.crc32(mv)
        try:
            for r in range(a.reps):
                t0 = time.perf_counter()
                crc, got = 0, 0
                with gt.open(path, "rb") as f:
                    if r == a.reps - 2 and a.reps > 2:          # one rep: read() of the whole file
                        got = len(f.read())
                    elif r == a.reps - 1 and a.reps > 1:          # last rep: readinto a reused buffer (no fresh pages)
                        box = bytearray(64 << 20)
                        while True:
                            k = f.readinto(box)
                            if not k:
                                break
                            got += k
                    else:
                        while True:
                            chunk = f.read(64 << 20)
                            if not chunk:
                                break
                            got += len(chunk)
                            if r == 0:
                                crc = zlib.crc32(chunk, crc)
                dt = time.perf_counter() - t0
                print(f"read rep{r}: {n / dt / 1e9:.2f} GB/s  ({dt * 1e3:.0f} ms)  bytes ok {got == n}" +
                      (f"  crc ok {crc == want} (checked inside the timed loop)" if r == 0 else "") +
                      ("  [readinto]" if r == a.reps - 1 and a.reps > 1 else "  [read() of everything]" if r == a.reps - 2 and a.reps > 2 else ""), flush=True)
        finally:
            os.unlink(path)
        return
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
