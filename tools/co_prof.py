import os, sys, time, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np, synthdata
from zlib_ng_b200 import zlib_ng
n = 512 << 20
base = synthdata.make("mixed", 256 << 20, 2)
data = np.tile(base, 2).tobytes()
zlib_ng.compress(data[:1 << 20])
def run():
    co = zlib_ng.compressobj(6, zlib_ng.DEFLATED, 31)
    tot = 0
    mv = memoryview(data)
    for off in range(0, n, 8 << 20):
        tot += len(co.compress(mv[off:off + (8 << 20)]))
    tot += len(co.flush())
    return tot
t = time.perf_counter(); run(); print("warm", time.perf_counter() - t)
pr = cProfile.Profile(); pr.enable(); t = time.perf_counter(); run(); dt = time.perf_counter() - t; pr.disable()
print(f"{n / dt / 1e9:.2f} GB/s")
pstats.Stats(pr).sort_stats("cumulative").print_stats(12)
