#!/usr/bin/env python3
"""Dev tool: host copy rate pageable -> pinned (b2d_host_copy), idle and while the GPU pipeline moves data."""
import ctypes, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np
from zlib_ng_b200 import _native as nat
L = nat.lib(); ctx = nat.Context(0, 64 << 20)
n = 2 << 30; piece = 64 << 20
src = np.random.randint(0, 255, n, dtype=np.uint8)
pin = ctx.host_alloc(4 * piece)
def run(threads, label):
    t0 = time.perf_counter()
    for i, off in enumerate(range(0, n, piece)):
        L.b2d_host_copy(pin + (i % 4) * piece, src.ctypes.data + off, piece, threads)
    dt = time.perf_counter() - t0
    print(f"{label} threads={threads}: {n/dt/1e9:.1f} GB/s  ({dt/ (n//piece)*1e3:.2f} ms per 64 MiB)", flush=True)
for t in (1, 4, 8, 12, 16):
    run(t, "idle")
# background: e2e compress from pinned host memory (H2D + kernels + D2H) on another context
ctx2 = nat.Context(0, 64 << 20)
hin = ctx2.host_alloc(1 << 30); cap = int(L.b2d_deflate_bound(1 << 30)); hout = ctx2.host_alloc(cap)
ctypes.memmove(hin, src.ctypes.data, 1 << 30)
stop = False
def bg():
    o, c = ctypes.c_uint64(0), ctypes.c_uint32(0)
    while not stop:
        L.b2d_compress_and_crc(ctx2.handle, 0, ctypes.c_void_p(hin), 1 << 30, None, 0, ctypes.c_void_p(hout), cap, ctypes.byref(o), ctypes.byref(c))
th = threading.Thread(target=bg); th.start(); time.sleep(0.5)
for t in (4, 8, 16):
    run(t, "with H2D+D2H busy (level 0)")
stop = True; th.join()
