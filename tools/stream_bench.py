#!/usr/bin/env python3
"""End-to-end throughput of gzip_ng_threaded.open(..., "wb"): the user-facing call, pageable host bytes in -> a gzip file,
this repo beside the reference on the host cores.  Runs in a FRESH process so that rep 0 is the cold number a
CLI-style "open, write, close" user sees (CUDA context, library load, first pinned buffers included).

    stream_bench.py --config 4 [--gib 16] [--devices 0,1,...]     BASELINE.json configs[4] (SURVEY §8d "Config 5"):
                        FASTQ-like seed 3 for the first half ‖ mixed-entropy seed 4 for the rest, level 6, threads = cores
    stream_bench.py --config 0                                    configs[0]: 256 MiB FASTQ-like (seed 1), level 1, threads=4

bench.py launches it (rank 0) and puts its one JSON line into `extra`.  Checks made inside: the file's trailer CRC32 /
ISIZE against the REFERENCE's crc32 of the input, the first --verify-mib MiB read back through the REFERENCE's reader
(gzip_ng_threaded.open(..., "rb") of oracle/_ref), and the whole file read back through this repo's GPU reader.
"""
import argparse
import io
import json
import os
import struct
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))
T_PROCESS = time.perf_counter()

import numpy as np  # noqa: E402
import synthdata  # noqa: E402

MIB, GIB = 1 << 20, 1 << 30


class NullSink(io.RawIOBase):
    def __init__(self):
        self.n = 0

    def writable(self):
        return True

    def write(self, b):
        self.n += len(b)
        return len(b)


def tiled(kind, n, seed, base_mib):
    base = synthdata.make(kind, min(n, base_mib * MIB), seed)
    return np.tile(base, (n + base.size - 1) // base.size)[:n]


def reference():
    pkg = os.path.join(ROOT, "oracle", "_ref", "pkg")
    if not os.path.isdir(os.path.join(pkg, "zlib_ng")):
        return None
    sys.path.insert(0, pkg)
    try:
        from zlib_ng import gzip_ng_threaded, zlib_ng  # type: ignore
        return zlib_ng, gzip_ng_threaded
    except Exception:
        return None


def write_pass(gt, sink, mv, n, step, level, threads, block_size):
    t0 = time.perf_counter()
    with gt.open(sink, "wb", compresslevel=level, threads=threads, block_size=block_size) as f:
        for off in range(0, n, step):
            f.write(mv[off:off + step])
    return time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=4, choices=[0, 4])
    ap.add_argument("--gib", type=float, default=None)
    ap.add_argument("--devices", default=None, help='B2D_DEVICES for this run, e.g. "0,1,2,3"')
    ap.add_argument("--threads", type=int, default=None)
    ap.add_argument("--write-mib", type=int, default=64)
    ap.add_argument("--base-mib", type=int, default=64)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--ref-sample-mib", type=int, default=512)
    ap.add_argument("--verify-mib", type=int, default=256)
    ap.add_argument("--no-ref", action="store_true")
    a = ap.parse_args()
    if a.devices:
        os.environ["B2D_DEVICES"] = a.devices
    ncores = len(os.sched_getaffinity(0))
    if a.config == 4:
        n = int((a.gib or 16.0) * GIB) // (256 * 1024) * (256 * 1024)
        level, threads, block = 6, a.threads or ncores, 1 << 20
        half = n // 2
        data = np.empty(n, dtype=np.uint8)
        data[:half] = tiled("fastq", half, 3, a.base_mib)
        data[half:] = tiled("mixed", n - half, 4, a.base_mib)
        what = f"{n / GIB:g} GiB: FASTQ-like (seed 3) for the first half ‖ mixed-entropy (seed 4), {a.base_mib} MiB segments tiled"
    else:
        n = int((a.gib or 0.25) * GIB)
        level, threads, block = 1, a.threads or 4, 1 << 20
        data = tiled("fastq", n, 1, max(a.base_mib, 256))
        what = f"{n / MIB:g} MiB FASTQ-like (seed 1)"
    mv = memoryview(data)
    step = a.write_mib * MIB
    res = {"config": a.config, "workload": what, "level": level, "threads": threads, "block_size": block, "host_cores": ncores,
           "bytes": n, "devices": os.environ.get("B2D_DEVICES", "all")}
    t_data = time.perf_counter()

    # ---- this repo, cold: first use of the library in this process
    from zlib_ng_b200 import _native as nat
    from zlib_ng_b200 import gzip_ng_threaded as ours
    shm = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    path = os.path.join(shm, f"b2d_stream_bench_{os.getpid()}.gz")
    times = []
    try:
        t_lib = time.perf_counter()
        dt = write_pass(ours, NullSink(), mv, n, step, level, threads, block)     # cold: the first use in this process
        times.append(dt)
        res["cold_s"] = dt
        res["cold_GBps"] = n / dt / 1e9
        res["cold_includes"] = "CUDA initialisation, contexts on every device used, first pinned buffers"
        res["n_gpus_used"] = min(threads, max(nat.lib().b2d_device_count(), 1)) if not a.devices else len(a.devices.split(","))
        for _ in range(max(a.reps - 1, 0)):
            times.append(write_pass(ours, NullSink(), mv, n, step, level, threads, block))
        warm = min(times[1:]) if len(times) > 1 else None
        res["warm_s"] = warm
        res["warm_GBps"] = n / warm / 1e9 if warm else None
        res["rep_s"] = times
        # one more pass into a real file (tmpfs), for the checks below; its time includes the page faults of the file
        with open(path, "wb") as fh:
            res["to_tmpfs_file_s"] = write_pass(ours, fh, mv, n, step, level, threads, block)
        res["file_ratio"] = os.path.getsize(path) / n
        del t_lib

        # ---- checks on that file
        ref = None if a.no_ref else reference()
        checks = {}
        blob_tail = None
        with open(path, "rb") as fh:
            fh.seek(0, os.SEEK_END)
            size = fh.tell()
            fh.seek(max(0, size - (1 << 20)))
            blob_tail = fh.read()
        import zlib
        crc_mod = ref[0] if ref else zlib
        t0 = time.perf_counter()
        want_crc = 0
        for off in range(0, n, 1 << 30):                      # >4 GiB inputs: chained, as the reference's own loop does
            want_crc = crc_mod.crc32(mv[off:off + (1 << 30)], want_crc)
        checks["crc_of_input_by"] = "reference zlib_ng.crc32" if ref else "stdlib zlib"
        # the data member's trailer sits right before the reference-style empty member (and the index members)
        from zlib_ng_b200.gzip_ng_threaded import _LOCATOR, _LOCATOR_MEMBER, _parse_extra_member
        got = _parse_extra_member(blob_tail[-_LOCATOR_MEMBER:], 0)
        if got:
            index_off = _LOCATOR.unpack(got[0])[4]
            with open(path, "rb") as fh:
                fh.seek(index_off - 20 - 8)
                crc_file, isize = struct.unpack("<II", fh.read(8))
            checks["trailer_crc_ok"] = crc_file == want_crc
            checks["trailer_isize_ok"] = isize == (n & 0xFFFFFFFF)
        t0 = time.perf_counter()
        back_crc, back_n = 0, 0
        with ours.open(path, "rb") as f:
            gpu_reader = f.raw._gpu is not None
            box = bytearray(64 * MIB)
            while True:
                k = f.readinto(box)
                if not k:
                    break
                back_crc = zlib.crc32(memoryview(box)[:k], back_crc)
                back_n += k
        checks["read_back_ok"] = back_n == n and back_crc == want_crc
        checks["read_back_on_gpu"] = gpu_reader
        checks["read_back_GBps_incl_host_crc"] = n / (time.perf_counter() - t0) / 1e9
        if ref:
            vm = min(a.verify_mib * MIB, n)
            with ref[1].open(path, "rb") as f:
                head = f.read(vm)
            checks["reference_reader_sample_ok"] = head == data[:vm].tobytes()
            checks["reference_reader_sample_MiB"] = vm // MIB
        res["checks"] = checks
        res["verify_ok"] = all(v for k, v in checks.items() if k.endswith("_ok"))
    finally:
        if os.path.exists(path):
            os.unlink(path)

    # ---- the reference on the host cores, same call, bounded sample of the same stream (both halves for config 4)
    ref = None if a.no_ref else reference()
    if ref:
        sn = min(a.ref_sample_mib * MIB, n)
        if a.config == 4:
            sample = np.concatenate([data[:sn // 2], data[n // 2:n // 2 + sn // 2]])
        else:
            sample = data[:sn]
        smv = memoryview(np.ascontiguousarray(sample))
        sink = NullSink()
        dt = write_pass(ref[1], sink, smv, sample.size, step, level, threads, block)
        res["reference"] = {"GBps": sample.size / dt / 1e9, "threads": threads, "cores": ncores, "ratio": sink.n / sample.size,
                            "sample": f"{sample.size // MIB} MiB of the same stream" + (" (half from each part)" if a.config == 4 else "")}
        sink2 = NullSink()
        write_pass(ours, sink2, smv, sample.size, step, level, threads, block)
        res["ratio_ours_on_reference_sample"] = sink2.n / sample.size
        res["size_ours_over_reference"] = sink2.n / sink.n
        res["speedup_warm"] = res["warm_GBps"] / res["reference"]["GBps"] if res.get("warm_GBps") else None
        res["speedup_cold"] = res["cold_GBps"] / res["reference"]["GBps"]
    res["data_gen_s"] = t_data - T_PROCESS
    print(json.dumps(res))


if __name__ == "__main__":
    main()
