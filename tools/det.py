import sys, os, zlib
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests'); sys.path.insert(0,'/root/repo/python-zlib-ng_b200')
import synthdata
from zlib_ng_b200 import _native as nat
if os.environ.get('B2D_LIB'): nat.LIB_PATH = os.environ['B2D_LIB']
from test_gpu_pipeline import deflate_host
from helpers import twin_deflate
data = synthdata.make("mixed", 5_000_000, 33).tobytes()
tw = twin_deflate(data, 6)
for mb in (1 << 20, 1 << 20, 4 << 20, 4 << 20, 16 << 20):
    c = nat.Context(0, mb)
    o = deflate_host(nat, c, data, 6)[0]
    c.close()
    ok = zlib.decompressobj(-15).decompress(o) == data
    # first difference vs twin
    d = next((i for i in range(min(len(o), len(tw))) if o[i] != tw[i]), None)
    print(mb >> 20, len(o), len(tw), "roundtrip", ok, "first diff vs twin", d)
