#!/usr/bin/env python3
"""The host-side ceiling of the end-to-end path: what this box can move with NO kernels running.

    python tools/copy_ceiling.py [max_devices] > profiles/copy_ceiling_r2.json

For N = 1, 2, 4, 8 GPUs at once: per GPU one thread copying 4 GiB pinned host -> device in 256 MiB pieces and one
thread copying 2.1 GiB device -> pinned host (the byte volumes of one bench step per GPU), concurrently, best of 3;
H2D alone and D2H alone as well.  Then memcpy pageable -> pinned at 1 / 4 / 8 / 16 / 32 threads (b2d_host_copy), the
copy the streaming writer makes of its caller's bytes."""
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
import numpy as np  # noqa: E402
from zlib_ng_b200 import _native as nat  # noqa: E402

MIB, GIB = 1 << 20, 1 << 30
L = nat.lib()
maxdev = min(int(sys.argv[1]) if len(sys.argv) > 1 else 8, max(L.b2d_device_count(), 1))
H2D, D2H, PIECE = 4 * GIB, int(2.1 * GIB) // MIB * MIB, 256 * MIB
res = {"h2d_bytes_per_gpu": H2D, "d2h_bytes_per_gpu": D2H, "piece": PIECE, "runs": []}

# two contexts per device: one stream each, so the two directions overlap
ctxs = [(nat.Context(d, 1 * MIB), nat.Context(d, 1 * MIB)) for d in range(maxdev)]
bufs = []
for a, b in ctxs:
    bufs.append({"h_in": a.host_alloc(PIECE * 2), "h_out": b.host_alloc(PIECE * 2), "d_in": a.device_alloc(PIECE * 2),
                 "d_out": b.device_alloc(PIECE * 2)})
    ctypes.memset(bufs[-1]["h_in"], 7, PIECE * 2)


def mover(ctx, kind, buf, total):
    done = 0
    i = 0
    while done < total:
        k = min(PIECE, total - done)
        o = (i & 1) * PIECE
        if kind == "h2d":
            ctx.h2d(buf["d_in"] + o, ctypes.c_void_p(buf["h_in"] + o), k)
        else:
            ctx.d2h(ctypes.c_void_p(buf["h_out"] + o), buf["d_out"] + o, k)
        done += k
        i += 1


def run(ndev, do_h2d, do_d2h):
    best = 1e9
    for _ in range(3):
        ths = []
        for d in range(ndev):
            if do_h2d:
                ths.append(threading.Thread(target=mover, args=(ctxs[d][0], "h2d", bufs[d], H2D)))
            if do_d2h:
                ths.append(threading.Thread(target=mover, args=(ctxs[d][1], "d2h", bufs[d], D2H)))
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        best = min(best, time.perf_counter() - t0)
    return best


n = 1
while n <= maxdev:
    t_both, t_in, t_out = run(n, True, True), run(n, True, False), run(n, False, True)
    res["runs"].append({"n_gpus": n, "both_s": t_both,
                        "both_input_GBps": n * H2D / t_both / 1e9,          # comparable with bench e2e (input bytes / time)
                        "both_dma_GBps": n * (H2D + D2H) / t_both / 1e9,
                        "h2d_only_GBps": n * H2D / t_in / 1e9, "d2h_only_GBps": n * D2H / t_out / 1e9})
    n *= 2

# pageable -> pinned memcpy, the streaming writer's one host copy
src = np.full(1 * GIB, 3, dtype=np.uint8)
dst = ctxs[0][0].host_alloc(1 * GIB)
res["memcpy_pageable_to_pinned_GBps"] = {}
for th in (1, 4, 8, 16, 32):
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        L.b2d_host_copy(ctypes.c_void_p(dst), ctypes.c_void_p(src.ctypes.data), src.size, th)
        best = min(best, time.perf_counter() - t0)
    res["memcpy_pageable_to_pinned_GBps"][str(th)] = src.size / best / 1e9
# the same copy while all GPUs move data (what the writer sees)
stop = False


def bg(d):
    while not stop:
        mover(ctxs[d][0], "h2d", bufs[d], PIECE * 4)


ths = [threading.Thread(target=bg, args=(d,)) for d in range(maxdev)]
for t in ths:
    t.start()
best = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    L.b2d_host_copy(ctypes.c_void_p(dst), ctypes.c_void_p(src.ctypes.data), src.size, 16)
    best = min(best, time.perf_counter() - t0)
stop = True
for t in ths:
    t.join()
res["memcpy_16_threads_while_all_gpus_h2d_GBps"] = src.size / best / 1e9
res["host_cpus"] = len(os.sched_getaffinity(0))
try:
    res["numa_nodes"] = len([x for x in os.listdir("/sys/devices/system/node") if x.startswith("node")])
except Exception:
    pass
print(json.dumps(res))
