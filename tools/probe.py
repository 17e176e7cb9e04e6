#!/usr/bin/env python3
"""Dev probe (not product): device-resident timing of the pipeline and checksums.

usage: probe.py [--kind fastq|mixed] [--mib N] [--levels 1,6] [--reps 3] [--batch-mib 256] [--crc-mib N]
"""
import argparse
import ctypes
import os
import sys
import time
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))

import numpy as np  # noqa: E402
import synthdata  # noqa: E402
from zlib_ng_b200 import _native as nat  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", default="fastq")
    ap.add_argument("--mib", type=int, default=256)
    ap.add_argument("--levels", default="1,6")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--batch-mib", type=int, default=256)
    ap.add_argument("--crc-mib", type=int, default=0)
    ap.add_argument("--verify", action="store_true")
    ap.add_argument("--inflate", action="store_true", help="also time b2d_inflate_device on each level's output")
    a = ap.parse_args()
    n = a.mib << 20
    t = time.time()
    if a.kind.startswith("seg"):       # one segment kind of the mixed generator: seg0 text .. seg5 base64
        base = synthdata._segment(int(a.kind[3:]), min(n, 64 << 20), synthdata._rng(7))
    else:
        base = synthdata.make(a.kind, min(n, 64 << 20), 1)
    data = np.tile(base, (n + base.size - 1) // base.size)[:n]
    print(f"gen {time.time() - t:.1f}s", flush=True)
    L = nat.lib()
    ctx = nat.Context(0, a.batch_mib << 20)
    d_in = ctx.device_alloc(n + 64)
    cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
    d_out = ctx.device_alloc(cap)
    ctx.h2d(d_in, data.ctypes.data_as(ctypes.c_void_p), n)
    for lvl in [int(x) for x in a.levels.split(",") if x]:
        for r in range(a.reps + 1):
            if r == a.reps:      # one more pass on a single stream with per-kernel events
                nat.check(L.b2d_profile(ctx.handle, 2))
                L.b2d_profile_read(ctx.handle, None, None, 1)
            out_n = ctypes.c_uint64(0)
            ms = ctypes.c_float(0)
            nat.check(L.b2d_deflate_device(ctx.handle, lvl, ctypes.c_void_p(d_in), 0, n, ctypes.c_void_p(d_out), cap,
                                           ctypes.byref(out_n), None, ctypes.byref(ms)))
            print(f"L{lvl} rep{r}: {ms.value:.2f} ms  {n / ms.value / 1e6:.2f} GB/s  ratio {out_n.value / n:.4f}", flush=True)
        sm = (ctypes.c_double * 6)()
        L.b2d_profile_read(ctx.handle, sm, None, 1)
        nat.check(L.b2d_profile(ctx.handle, 0))
        print("   stage ms alone:", " ".join(f"{nm}={sm[i]:.2f}" for i, nm in enumerate(["chains", "match", "parse", "plan", "scan", "emit"])), flush=True)
        if a.inflate:
            cnt = ctypes.c_uint64(0)
            nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
            offs = (ctypes.c_uint64 * cnt.value)()
            nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
            d_back = ctx.device_alloc(n + 64)
            for r in range(a.reps):
                ims = ctypes.c_float(0)
                nat.check(L.b2d_inflate_device(ctx.handle, ctypes.c_void_p(d_out), out_n.value, offs, cnt.value - 1, None, 0,
                                               ctypes.c_void_p(d_back), n, ctypes.byref(ims)))
                print(f"   inflate rep{r}: {ims.value:.2f} ms  {n / ims.value / 1e6:.2f} GB/s (output bytes)", flush=True)
            v1, v2 = ctypes.c_uint32(0), ctypes.c_uint32(0)
            nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_back), n, ctypes.byref(v1), None))
            nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_in), n, ctypes.byref(v2), None))
            print("   inflate crc matches input:", v1.value == v2.value, flush=True)
            ctx.device_free(d_back)
        if a.verify:
            host = np.empty(out_n.value, dtype=np.uint8)
            ctx.d2h(host.ctypes.data_as(ctypes.c_void_p), d_out, out_n.value)
            back = zlib.decompressobj(-15).decompress(host.tobytes())
            print("round-trip", back == data.tobytes())
    if a.crc_mib:
        m = min(a.crc_mib << 20, n)
        for name, fn in (("crc32", L.b2d_crc32_device), ("adler32", L.b2d_adler32_device)):
            for r in range(a.reps):
                v = ctypes.c_uint32(0)
                ms = ctypes.c_float(0)
                nat.check(fn(ctx.handle, 0 if name == "crc32" else 1, ctypes.c_void_p(d_in), m, ctypes.byref(v), ctypes.byref(ms)))
                print(f"{name} rep{r}: {ms.value:.3f} ms  {m / ms.value / 1e6:.1f} GB/s  value {v.value:#x}", flush=True)
        ref = zlib.crc32(data[:m]) if m <= (1 << 30) else None
        print("crc ref", None if ref is None else hex(ref))


if __name__ == "__main__":
    main()
