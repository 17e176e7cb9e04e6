#!/usr/bin/env python3
"""Dev tool: gzip_ng_threaded.open(..., "rb") GB/s (of uncompressed bytes) on a file this writer indexed and on a BGZF
file of the same data, readinto a reused buffer / read(64 MiB) / read(), beside the reference's reader on the same files.
usage: reader_bench.py [gib]"""
import os
import struct
import sys
import time
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np  # noqa: E402
import synthdata  # noqa: E402
from zlib_ng_b200 import gzip_ng_threaded as ours  # noqa: E402

n = int(float(sys.argv[1]) * (1 << 30)) if len(sys.argv) > 1 else 4 << 30
base = synthdata.make("mixed", 64 << 20, 2)
data = np.tile(base, n // base.size)
mv = memoryview(data)
shm = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
p1, p2 = os.path.join(shm, "b2d_rb_indexed.gz"), os.path.join(shm, "b2d_rb.bgzf.gz")
with ours.open(p1, "wb", threads=16) as f:
    f.write(mv)
nb = min(n, 1 << 30)                       # the BGZF file is written by CPython (slow): 1 GiB
with open(p2, "wb") as f:
    for o in range(0, nb, 65280):
        piece = bytes(mv[o:min(o + 65280, nb)])
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = c.compress(piece) + c.flush()
        f.write(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", len(body) + 25) + body
                + struct.pack("<II", zlib.crc32(piece), len(piece)))


def run(mod, path, total, how):
    t0 = time.perf_counter()
    got = 0
    with mod.open(path, "rb") as f:
        if how == "readinto":
            box = bytearray(64 << 20)
            while True:
                k = f.readinto(box)
                if not k:
                    break
                got += k
        elif how == "read64":
            while True:
                b = f.read(64 << 20)
                if not b:
                    break
                got += len(b)
        else:
            got = len(f.read())
    assert got == total, (got, total)
    return total / (time.perf_counter() - t0) / 1e9


try:
    for name, path, total in (("indexed", p1, n), ("bgzf", p2, nb)):
        for how in ("readinto", "read64", "read"):
            best = max(run(ours, path, total, how) for _ in range(3))
            print(f"ours {name} {how}: {best:.2f} GB/s", flush=True)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "pkg"))
    try:
        from zlib_ng import gzip_ng_threaded as ref  # type: ignore
        for name, path, total in (("indexed", p1, n), ("bgzf", p2, nb)):
            if total > (1 << 30):
                continue
            print(f"reference {name} readinto: {run(ref, path, total, 'readinto'):.2f} GB/s", flush=True)
    except ImportError:
        pass
finally:
    os.unlink(p1)
    os.unlink(p2)
