#!/usr/bin/env python3
"""Dev tool: b2d_compress_and_crc_multi alone (pinned host in -> pinned host out) on 1..N GPUs of this box, per batch size.
usage: multi_e2e.py [max_devices] [gib_per_device]     (B2D_TRACE=1 prints where each device thread spent its time)"""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np  # noqa: E402
import synthdata  # noqa: E402
from zlib_ng_b200 import _native as nat  # noqa: E402

L = nat.lib()
maxdev = min(int(sys.argv[1]) if len(sys.argv) > 1 else 8, L.b2d_device_count())
gib = float(sys.argv[2]) if len(sys.argv) > 2 else 4.0
base = synthdata.make("mixed", 64 << 20, 2)
nd = 1
while nd <= maxdev:
    n = int(gib * (1 << 30)) * nd
    cap = int(L.b2d_deflate_bound(n)) + 64
    h_in, h_out = nat.pinned_alloc(n), nat.pinned_alloc(cap)
    hv = np.frombuffer((ctypes.c_uint8 * n).from_address(h_in), dtype=np.uint8)
    for off in range(0, n, base.size):
        hv[off:off + base.size] = base[:min(base.size, n - off)]
    for batch in (64, 256):
        m = nat.MultiContext(list(range(nd)), batch << 20)
        m.compress_and_crc(6, h_in, n, None, 0, h_out, cap)
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            m.compress_and_crc(6, h_in, n, None, 0, h_out, cap)
            best = min(best, time.perf_counter() - t0)
        print(f"devices={nd} batch_mib={batch}: {n / best / 1e9:.1f} GB/s ({best * 1e3:.0f} ms for {n >> 30} GiB)", flush=True)
        m.close()
    del hv
    nat.pinned_free(h_in)
    nat.pinned_free(h_out)
    nd *= 2
