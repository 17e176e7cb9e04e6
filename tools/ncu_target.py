#!/usr/bin/env python3
"""Target for the ncu --set full capture: twelve device-resident passes (level 6, or B2D_NCU_LEVEL) over one 256 MiB batch of
the bench workload, then inflates of the result (tools/run_ncu_r2.sh skips the first launches and captures one of each kernel)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import synthdata  # noqa: E402
from zlib_ng_b200 import _native as nat  # noqa: E402

L = nat.lib()
LEVEL = int(os.environ.get("B2D_NCU_LEVEL", "6"))
n = 256 << 20
base = synthdata.make("mixed", n, 2)
ctx = nat.Context(0, n)
d_in = ctx.device_alloc(n + 64)
cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
d_out = ctx.device_alloc(cap)
d_back = ctx.device_alloc(n + 64)
ctx.h2d(d_in, base.ctypes.data_as(ctypes.c_void_p), n)
out_n, crc = ctypes.c_uint64(0), ctypes.c_uint32(0)
for _ in range(12):
    nat.check(L.b2d_deflate_device(ctx.handle, LEVEL, ctypes.c_void_p(d_in), 0, n, ctypes.c_void_p(d_out), cap, ctypes.byref(out_n), None, None))
cnt = ctypes.c_uint64(0)
nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
offs = (ctypes.c_uint64 * cnt.value)()
nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
for _ in range(2):
    nat.check(L.b2d_inflate_device(ctx.handle, ctypes.c_void_p(d_out), out_n.value, offs, cnt.value - 1, None, 0, ctypes.c_void_p(d_back), n, None))
v = ctypes.c_uint32(0)
nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_back), n, ctypes.byref(v), None))
w = ctypes.c_uint32(0)
nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_in), n, ctypes.byref(w), None))
assert v.value == w.value, "inflate under the profiler gave different bytes"
print("ok", out_n.value / n)
