#!/usr/bin/env python3
"""Dev tool: where does gzip_ng_threaded.open(..., 'wb') spend its wall time?  Wraps the writer's steps with timers."""
import io, os, sys, time, threading, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np
import synthdata
from zlib_ng_b200 import gzip_ng_threaded as gt

T = collections.defaultdict(float); C = collections.defaultdict(int); lock = threading.Lock()
def timed(name, fn):
    def w(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            dt = time.perf_counter() - t0
            with lock:
                T[name] += dt; C[name] += 1
    return w
gt._parallel_copy = timed("main.parallel_copy", gt._parallel_copy)
W = gt._ThreadedGzipWriter
W._take_buffer = timed("main.take_buffer", W._take_buffer)
W._submit_current = timed("main.submit(put)", W._submit_current)
W._run_job = timed("worker.run_job", W._run_job)
W._end_gzip_stream = timed("main.end_stream", W._end_gzip_stream)

class Null(io.RawIOBase):
    n = 0
    def writable(self): return True
    def write(self, b):
        self.n += len(b); return len(b)

n = int(float(sys.argv[1]) * (1 << 30)) if len(sys.argv) > 1 else 4 << 30
base = synthdata.make("mixed", 256 << 20, 2)
data = np.tile(base, (n + base.size - 1) // base.size)[:n]
mv = memoryview(data)
for rep in range(3):
    T.clear(); C.clear()
    s = Null(); t0 = time.perf_counter()
    with gt.open(s, "wb", compresslevel=6, threads=16) as f:
        for off in range(0, n, 64 << 20):
            f.write(mv[off:off + (64 << 20)])
    dt = time.perf_counter() - t0
    print(f"rep{rep}: {n/dt/1e9:.2f} GB/s {dt*1e3:.0f} ms")
    for k in sorted(T):
        print(f"   {k:22s} total {T[k]*1e3:8.1f} ms  calls {C[k]:4d}  avg {T[k]/max(C[k],1)*1e3:7.2f} ms")
