#!/usr/bin/env python3
"""Target for an ncu launch list of the inflate kernels: one device-resident inflate per sub-block of 64 MiB and of 512 MiB."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import synthdata  # noqa: E402
from zlib_ng_b200 import _native as nat  # noqa: E402

L = nat.lib()
N = 512 << 20
base = synthdata.make("mixed", 256 << 20, 2)
ctx = nat.Context(0, 256 << 20)
d_in = ctx.device_alloc(N + 64)
for o in range(0, N, base.size):
    ctx.h2d(d_in + o, base.ctypes.data_as(ctypes.c_void_p), base.size)
cap = (int(L.b2d_deflate_bound(N)) + 3) & ~3
d_out = ctx.device_alloc(cap)
d_back = ctx.device_alloc(N + 64)
for mib in (64, 512):
    n = mib << 20
    out_n = ctypes.c_uint64(0)
    nat.check(L.b2d_deflate_device(ctx.handle, 6, ctypes.c_void_p(d_in), 0, n, ctypes.c_void_p(d_out), cap, ctypes.byref(out_n), None, None))
    cnt = ctypes.c_uint64(0)
    nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
    offs = (ctypes.c_uint64 * cnt.value)()
    nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
    sc = ctypes.c_uint64(0)
    nat.check(L.b2d_subblock_bits(ctx.handle, None, 0, ctypes.byref(sc)))
    subs = (ctypes.c_uint32 * sc.value)()
    nat.check(L.b2d_subblock_bits(ctx.handle, subs, sc.value, ctypes.byref(sc)))
    for _ in range(2):
        nat.check(L.b2d_inflate_device_sub(ctx.handle, ctypes.c_void_p(d_out), out_n.value, offs, subs, cnt.value - 1, None, 0,
                                           ctypes.c_void_p(d_back), n, None))
print("ok")
