#!/usr/bin/env python3
"""Dev tool: gzip_ng_threaded.open(..., "wb") GB/s against B2D_STREAM_DEPTH / B2D_STREAM_BATCH_MIB (each combination in its own
process), pageable 64 MiB writes, one pageable write and one write from a pinned_buffer.  usage: writer_sweep.py [gib] [devices]"""
import io
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
    import numpy as np
    import synthdata
    from zlib_ng_b200 import gzip_ng_threaded as gzt
    n = int(float(sys.argv[2]) * (1 << 30))
    base = synthdata.make("mixed", 64 << 20, 2)
    data = np.tile(base, n // base.size)
    mv = memoryview(data)

    class Null(io.RawIOBase):
        n = 0

        def writable(self):
            return True

        def write(self, b):
            self.n += len(b)
            return len(b)
    ncores = len(os.sched_getaffinity(0))
    res = {}

    def best_of(fn, reps):
        best = 0
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            best = max(best, n / (time.perf_counter() - t0) / 1e9)
        return round(best, 2)

    def chunks():
        with gzt.open(Null(), "wb", compresslevel=6, threads=ncores) as f:
            for off in range(0, n, 64 << 20):
                f.write(mv[off:off + (64 << 20)])

    def one():
        with gzt.open(Null(), "wb", compresslevel=6, threads=ncores) as f:
            f.write(mv)
    res["pageable_64MiB_writes"] = best_of(chunks, 4)
    res["pageable_one_write"] = best_of(one, 3)
    with gzt.pinned_buffer(n) as pb:
        np.frombuffer(pb.view, dtype=np.uint8)[:] = data

        def pinned():
            with gzt.open(Null(), "wb", compresslevel=6, threads=ncores) as f:
                f.write(pb.view)
        res["pinned_one_write"] = best_of(pinned, 3)
    print(json.dumps(res))
    sys.exit(0)

gib = sys.argv[1] if len(sys.argv) > 1 else "4"
devs = sys.argv[2] if len(sys.argv) > 2 else "0"
for depth, batch in ((2, 256), (3, 128), (3, 256), (4, 128), (2, 512)):
    env = dict(os.environ, B2D_STREAM_DEPTH=str(depth), B2D_STREAM_BATCH_MIB=str(batch), B2D_DEVICES=devs)
    p = subprocess.run([sys.executable, __file__, "--child", gib], env=env, capture_output=True, text=True)
    print(f"depth={depth} batch_mib={batch} devices={devs}", p.stdout.strip() or p.stderr[-300:], flush=True)
