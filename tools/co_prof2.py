import cProfile, pstats, sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np, synthdata
from zlib_ng_b200 import zlib_ng
n = 1 << 30
base = synthdata.make("mixed", 64 << 20, 2)
data = np.tile(base, n // base.size).tobytes()
mv = memoryview(data)
def run():
    co = zlib_ng.compressobj(6, zlib_ng.DEFLATED, 31)
    tot = 0
    for off in range(0, n, 8 << 20):
        tot += len(co.compress(mv[off:off + (8 << 20)]))
    tot += len(co.flush())
    return tot
run()
t = time.perf_counter(); run(); print("GB/s", n / (time.perf_counter() - t) / 1e9)
cProfile.run("run()", "/tmp/co.prof")
pstats.Stats("/tmp/co.prof").sort_stats("tottime").print_stats(12)
