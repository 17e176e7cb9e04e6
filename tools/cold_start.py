#!/usr/bin/env python3
"""Where the first call of a process spends its time.  Run in a FRESH process:
    python tools/cold_start.py [n_devices]
Prints one JSON line: seconds for loading the library, creating the contexts, the first pinned buffers,
the first compress call, and a second (warm) call of the same size."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
t_start = time.perf_counter()
import numpy as np  # noqa: E402

ndev = int(sys.argv[1]) if len(sys.argv) > 1 else 1
MIB = 1 << 20
out = {}
t0 = time.perf_counter()
from zlib_ng_b200 import _native as nat  # noqa: E402
L = nat.lib()
out["load_lib_s"] = time.perf_counter() - t0
t0 = time.perf_counter()
ndev = min(ndev, max(L.b2d_device_count(), 1))
out["device_count_s"] = time.perf_counter() - t0
t0 = time.perf_counter()
m = nat.MultiContext(list(range(ndev)), 64 * MIB)
out["mctx_create_s"] = time.perf_counter() - t0
n = 256 * MIB * ndev
cap = int(L.b2d_deflate_bound(n)) + 64
rng = np.random.default_rng(1)
data = np.tile(rng.integers(0, 64, 1 * MIB, dtype=np.uint8), n // MIB)
one = nat.Context(0, 1 * MIB)
t0 = time.perf_counter()
h_in, h_out = one.host_alloc(n), one.host_alloc(cap)
out["pinned_alloc_s"] = time.perf_counter() - t0
out["pinned_alloc_GiB"] = (n + cap) / 2**30
ctypes.memmove(h_in, data.ctypes.data, n)
for name in ("first_call_s", "second_call_s", "third_call_s"):
    t0 = time.perf_counter()
    m.compress_and_crc(6, h_in, n, None, 0, h_out, cap)
    out[name] = time.perf_counter() - t0
out["call_GiB"] = n / 2**30
out["n_devices"] = ndev
out["total_s"] = time.perf_counter() - t_start
print(json.dumps(out))
