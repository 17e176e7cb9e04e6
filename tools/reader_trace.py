import os, sys, time
ROOT='/root/repo'
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np, synthdata
from zlib_ng_b200 import gzip_ng_threaded as ours
n = 4 << 30
base = synthdata.make("mixed", 64 << 20, 2)
data = np.tile(base, n // base.size); mv = memoryview(data)
p1 = "/dev/shm/b2d_rb_indexed.gz"
with ours.open(p1, "wb", threads=16) as f: f.write(mv)
box = bytearray(64 << 20)
for rep in range(3):
    t0 = time.perf_counter(); got = 0
    with ours.open(p1, "rb") as f:
        t1 = time.perf_counter()
        while True:
            k = f.readinto(box)
            if not k: break
            got += k
        t2 = time.perf_counter()
    t3 = time.perf_counter()
    print("rep", rep, "open %.1f ms, loop %.1f ms, close %.1f ms -> %.2f GB/s" % (1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), n/(t3-t0)/1e9), flush=True)
os.unlink(p1)
