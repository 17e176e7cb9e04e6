#!/usr/bin/env python3
"""bench.py — headline benchmark: gzip/DEFLATE compress throughput (input GB/s).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle/_ref)

Workload (BASELINE.json configs[3]): level-6 compression of 4 GiB of
mixed-entropy data per GPU, 128 KiB device blocks primed with the previous
32 KiB; one "step" = one pass of the whole pipeline over those 4 GiB.  Ranks are
independent shards (contiguous ranges, no collective on the data path), so
scaling is "weak"; torch.distributed (NCCL) is used only for the barrier and the
max-over-ranks of the device times.

value   device-resident: input already in HBM, output left in HBM, CUDA events
e2e     the same pass through the C ABI with pinned HOST buffers
        (b2d_deflate_host: H2D + kernels + CRC + D2H inside the timed region)
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "python-zlib-ng_b200"))

import numpy as np  # noqa: E402

import synthdata  # noqa: E402

GIB = 1 << 30
MIB = 1 << 20
METRIC = "gzip_compress_input_GBps_L6"
STAGES = ["k_chains", "k_match", "k_parse", "k_plan", "k_scan", "k_emit"]
# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per INPUT byte of a launch, from the one
# `ncu --set full` capture of a 256 MiB level-6 batch of this workload (profiles/ncu_r1_e_summary.txt §2)
NCU_DRAM_BYTES_PER_INPUT_BYTE = {"k_chains": 3.57, "k_match": 7.33, "k_parse": 9.32, "k_plan": 0.063,
                                 "k_scan": 0.004, "k_emit": 2.01}


def make_workload(kind, n, seed, base_mib):
    """`n` bytes: a seeded `base_mib` MiB segment of the synthetic corpus, tiled.  The tile period is far
    beyond DEFLATE's 32 KiB window, so tiling does not change the work per byte or the ratio."""
    base = synthdata.make(kind, min(n, base_mib * MIB), seed)
    reps = (n + base.size - 1) // base.size
    return base, reps


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons of one GPU every 100 ms while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        self.ready = threading.Event()      # NVML is up and the first sample is in: the timed region may start

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                     0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                self.ready.set()
                time.sleep(0.02)
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"sampler_error:{type(e).__name__}")
            self.ready.set()

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        # the first sample predates the timed region (taken while waiting for `ready`)
        under_load = self.samples[1:] if len(self.samples) > 1 else self.samples
        med = float(np.median(under_load)) if under_load else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


_JSON_FD = None


def claim_stdout():
    """Keep stdout for the one JSON line: anything else (NCCL's version banner, library chatter written
    to fd 1 from C) goes to stderr."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        os.write(1, data)
    else:
        os.write(_JSON_FD, data)


def bind_to_gpu_numa_node(local):
    """Run this rank's host threads (and therefore first-touch its pinned buffers) on the NUMA node
    the GPU hangs off: with 4-8 ranks the end-to-end number is otherwise limited by the inter-socket
    link, not by PCIe.  Returns a short description for the config, or None when nothing was done."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:      # nvml pads the domain to 8 hex digits, sysfs uses 4
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return f"numa node {node} ({len(cpus)} cpus)"
    except Exception:
        return None


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    return rank, world, local, dist


def barrier(dist, local):
    if dist is not None:
        import torch
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize()


def allreduce_max(dist, local, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allreduce_sum(dist, local, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def reference_module():
    pkg = os.path.join(ROOT, "oracle", "_ref", "pkg")
    if os.path.isdir(os.path.join(pkg, "zlib_ng")):
        sys.path.insert(0, pkg)
        try:
            from zlib_ng import gzip_ng_threaded, zlib_ng  # type: ignore
            return zlib_ng, gzip_ng_threaded
        except Exception:
            return None
    return None


class NullSink:
    def __init__(self):
        self.n = 0

    def write(self, b):
        self.n += len(b)
        return len(b)

    def flush(self):
        pass


def cpu_reference_pass(sample: np.ndarray, level: int, threads: int):
    """The reference's own hot path on the host cores: gzip_ng_threaded.open('wb', threads=all).
    Falls back to the repo's CPU twin (a port of OUR pipeline, single thread) if oracle/_ref is absent."""
    ref = reference_module()
    mv = memoryview(sample)
    if ref is not None:
        _, gzt = ref
        sink = NullSink()
        t0 = time.perf_counter()
        with gzt.open(sink, "wb", compresslevel=level, threads=threads) as f:
            for i in range(0, sample.size, 8 * MIB):
                f.write(mv[i:i + 8 * MIB])
        dt = time.perf_counter() - t0
        return dt, sink.n, "reference", threads
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import twin_deflate
    t0 = time.perf_counter()
    out = twin_deflate(sample.tobytes(), level)
    return time.perf_counter() - t0, len(out), "port", 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gib", type=float, default=4.0, help="input GiB per GPU")
    ap.add_argument("--level", type=int, default=6)
    ap.add_argument("--kind", default="mixed")
    ap.add_argument("--batch-mib", type=int, default=256)
    ap.add_argument("--base-mib", type=int, default=256)
    ap.add_argument("--cpu-sample-mib", type=int, default=512)
    ap.add_argument("--no-extras", action="store_true")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    n = int(a.gib * GIB) // (128 * 1024) * (128 * 1024)
    ncores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    config = {"workload": f"level-{a.level} DEFLATE of {a.gib:g} GiB {a.kind}-entropy synthetic data per GPU, "
                          f"128 KiB blocks + 32 KiB dictionary (BASELINE.json configs[3])",
              "bytes_per_gpu": n, "level": a.level, "block": 131072, "batch_mib": a.batch_mib,
              "l2": "inputs (GiB) far larger than the 126 MB L2; no flush needed",
              "parallelism": f"{a.gpus} independent shard(s), no collective"}

    if a.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        if rank != 0:
            return 0
        sample_n = min(n, a.cpu_sample_mib * MIB)
        base, _ = make_workload(a.kind, sample_n, 2, a.base_mib)
        sample = np.ascontiguousarray(np.tile(base, (sample_n + base.size - 1) // base.size)[:sample_n])
        for _ in range(a.warmup):
            cpu_reference_pass(sample[:min(sample_n, 64 * MIB)], a.level, ncores)
        t_total, kind, cores, out_n = 0.0, "reference", ncores, 0
        for _ in range(a.steps):
            dt, out_n, kind, cores = cpu_reference_pass(sample, a.level, ncores)
            t_total += dt
        val = sample_n * a.steps / t_total / 1e9
        line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": a.gpus, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": 1e3 * t_total / a.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": kind,
                                 "sample": f"first {sample_n // MIB} MiB of the workload per step through "
                                           f"gzip_ng_threaded.open('wb', compresslevel={a.level}, threads={cores}), "
                                           f"ratio {out_n / sample_n:.4f}"},
                "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit_line(line)
        return 0

    # ------------------------------------------------------------------ this repo's arm
    claim_stdout()
    rank, world, local, dist = dist_setup(a.gpus)
    config["host_binding"] = bind_to_gpu_numa_node(local) if world > 1 else None
    # synthetic data first (numpy), CUDA afterwards
    base, reps = make_workload(a.kind, n, 2 + rank, a.base_mib)
    from zlib_ng_b200 import _native as nat
    L = nat.lib()
    ctx = nat.Context(local, a.batch_mib * MIB)
    d_in = ctx.device_alloc(n + 64)
    cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
    d_out = ctx.device_alloc(cap)
    for r in range(reps):
        m = min(base.size, n - r * base.size)
        ctx.h2d(d_in + r * base.size, base.ctypes.data_as(ctypes.c_void_p), m)

    def device_pass(level=a.level, nbytes=n, want_crc=True):
        out_n, crc, ms = ctypes.c_uint64(0), ctypes.c_uint32(0), ctypes.c_float(0)
        nat.check(L.b2d_deflate_device(ctx.handle, level, ctypes.c_void_p(d_in), 0, nbytes, ctypes.c_void_p(d_out), cap,
                                       ctypes.byref(out_n), ctypes.byref(crc) if want_crc else None, ctypes.byref(ms)))
        return ms.value, out_n.value, crc.value

    for _ in range(a.warmup):
        device_pass()
    nat.check(L.b2d_profile(ctx.handle, 1))
    L.b2d_profile_read(ctx.handle, None, None, 1)
    launches0 = ctx.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    sampler.ready.wait(timeout=20)
    barrier(dist, local)
    ctx.sync()
    t_wall0 = time.perf_counter()
    dev_ms, out_n, crc = 0.0, 0, 0
    for _ in range(a.steps):
        ms, out_n, crc = device_pass()
        dev_ms += ms
    ctx.sync()
    barrier(dist, local)
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop()
    launches = ctx.launch_count() - launches0
    stage_ms = (ctypes.c_double * 6)()
    stage_n = (ctypes.c_uint64 * 6)()
    L.b2d_profile_read(ctx.handle, stage_ms, stage_n, 1)
    nat.check(L.b2d_profile(ctx.handle, 0))
    # the same kernels timed alone (one stream, no overlap between batches): explains the shares
    nat.check(L.b2d_profile(ctx.handle, 2))
    device_pass()
    alone_ms = (ctypes.c_double * 6)()
    L.b2d_profile_read(ctx.handle, alone_ms, None, 1)
    nat.check(L.b2d_profile(ctx.handle, 0))
    dev_ms_max = allreduce_max(dist, local, dev_ms)
    value = world * n * a.steps / (dev_ms_max * 1e-3) / 1e9
    ratio = out_n / n

    # roofline of the dominant kernel (largest share of the step), from the live per-kernel events
    shares = {STAGES[i]: stage_ms[i] for i in range(6)}
    top = max(range(6), key=lambda i: stage_ms[i])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes_per_launch = (1.0 + ratio) * n * a.steps / max(int(stage_n[top]), 1)
    achieved = alg_bytes_per_launch / (stage_ms[top] / max(int(stage_n[top]), 1) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": STAGES[top], "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak,
                "traffic": NCU_DRAM_BYTES_PER_INPUT_BYTE[STAGES[top]] * n * a.steps / max(int(stage_n[top]), 1),
                "traffic_source": "ncu --set full capture of one launch (profiles/ncu_r1_e_summary.txt), DRAM bytes per "
                                  "input byte x input bytes per launch",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
                "algorithmic_bytes_per_launch": alg_bytes_per_launch,
                "kernel_ms_per_step": {k: v / a.steps for k, v in shares.items()},
                "kernel_ms_per_step_alone": {STAGES[i]: alone_ms[i] for i in range(6)},
                "note": "DEFLATE is bound by shared-memory / integer work, not HBM; see DESIGN.md §5"}

    # ------------------------------------------------------------------ end to end through the C ABI, host buffers
    e2e = None
    stream_extra = {}
    try:
        h_in = ctx.host_alloc(n)
        h_out = ctx.host_alloc(cap)
        hv = np.frombuffer((ctypes.c_uint8 * n).from_address(h_in), dtype=np.uint8)
        for r in range(reps):
            m = min(base.size, n - r * base.size)
            hv[r * base.size:r * base.size + m] = base[:m]
        def host_pass():
            o, c = ctypes.c_uint64(0), ctypes.c_uint32(0)
            nat.check(L.b2d_compress_and_crc(ctx.handle, a.level, ctypes.c_void_p(h_in), n, None, 0,
                                             ctypes.c_void_p(h_out), cap, ctypes.byref(o), ctypes.byref(c)))
            return o.value, c.value
        host_pass()
        barrier(dist, local)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            o_n, c_e2e = host_pass()
        t_e2e = time.perf_counter() - t0
        t_e2e = allreduce_max(dist, local, t_e2e)
        assert c_e2e == crc and o_n == out_n, "host path disagrees with device-resident path"
        e2e = {"value": world * n * a.steps / t_e2e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": n,
               "d2h_bytes_per_step": o_n, "ms_per_step": 1e3 * t_e2e / a.steps,
               "api": "b2d_compress_and_crc(pinned host in -> pinned host out), the call behind "
                      "_ParallelCompress.compress_and_crc / gzip_ng_threaded"}
        if rank == 0 and world == 1 and not a.no_extras:
            # the user-facing call: gzip_ng_threaded.open(sink, "wb", threads=N).write(host bytes) -> gzip bytes
            import io
            from zlib_ng_b200 import gzip_ng_threaded as gzt

            class _Null(io.RawIOBase):
                n = 0

                def writable(self):
                    return True

                def write(self, b):
                    self.n += len(b)
                    return len(b)
            best = 0.0
            for _ in range(3):
                sink = _Null()
                t0 = time.perf_counter()
                with gzt.open(sink, "wb", compresslevel=a.level, threads=ncores) as f:
                    for off in range(0, n, 64 * MIB):
                        f.write(memoryview(hv[off:off + 64 * MIB]))
                best = max(best, n / (time.perf_counter() - t0) / 1e9)
            stream_extra = {"gzip_ng_threaded_open_wb_GBps": best, "gzip_ng_threaded_ratio": sink.n / n}
        ctx.host_free(h_in)
        ctx.host_free(h_out)
    except Exception as e:  # report, never fake
        e2e = {"value": None, "unit": "GB/s", "error": f"{type(e).__name__}: {e}"}

    # ------------------------------------------------------------------ extras + CPU baseline (rank 0, N == 1 only)
    extra, cpu = {}, None
    extra.update(stream_extra)
    if rank == 0 and world == 1 and not a.no_extras:
        # correctness of this very run: round-trip of the first 64 MiB + CRC of everything
        import zlib
        head_n = 64 * MIB
        _, o1, c1 = device_pass(nbytes=head_n)
        host = np.empty(o1, dtype=np.uint8)
        ctx.d2h(host.ctypes.data_as(ctypes.c_void_p), d_out, o1)
        want = np.tile(base, (head_n + base.size - 1) // base.size)[:head_n].tobytes()
        extra["roundtrip_64MiB_ok"] = zlib.decompressobj(-15).decompress(host.tobytes()) == want
        extra["crc_ok"] = c1 == zlib.crc32(want)
        # the other side of the wire: block-parallel inflate of this run's own output,
        # device-resident, checked with the CRC kernel; beside it the reference's inflate (one thread, as its
        # reader runs it) on the first 256 MiB worth of the same stream
        try:
            _, o_full, c_full = device_pass()
            cnt = ctypes.c_uint64(0)
            nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
            offs = (ctypes.c_uint64 * cnt.value)()
            nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
            d_back = ctx.device_alloc(n + 64)
            best = 1e9
            for _ in range(3):
                ims = ctypes.c_float(0)
                nat.check(L.b2d_inflate_device(ctx.handle, ctypes.c_void_p(d_out), o_full, offs, cnt.value - 1, None, 0,
                                               ctypes.c_void_p(d_back), n, ctypes.byref(ims)))
                best = min(best, ims.value)
            v = ctypes.c_uint32(0)
            nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_back), n, ctypes.byref(v), None))
            ctx.device_free(d_back)
            extra["inflate_GBps"] = n / best / 1e6
            extra["inflate_crc_ok"] = v.value == c_full
            ref = reference_module()
            nblk = min(2048, cnt.value - 1)
            comp = np.empty(offs[nblk], dtype=np.uint8)
            ctx.d2h(comp.ctypes.data_as(ctypes.c_void_p), d_out, int(offs[nblk]))
            zmod = ref[0] if ref is not None else zlib
            t0 = time.perf_counter()
            back_n = len(zmod.decompressobj(-15).decompress(comp.tobytes()))
            extra["inflate_reference_1thread_GBps"] = back_n / (time.perf_counter() - t0) / 1e9
            extra["inflate_reference_kind"] = "reference zlib_ng.decompressobj" if ref is not None else "stdlib zlib"
        except Exception as e:  # report, never fake
            extra["inflate_error"] = f"{type(e).__name__}: {e}"
        ms1 = min(device_pass(level=1)[0] for _ in range(3))
        extra["L1_mixed_GBps"] = n / ms1 / 1e6
        for name, fn, start in (("crc32", L.b2d_crc32_device, 0), ("adler32", L.b2d_adler32_device, 1)):
            best = 1e9
            for _ in range(4):
                v, ms = ctypes.c_uint32(0), ctypes.c_float(0)
                nat.check(fn(ctx.handle, start, ctypes.c_void_p(d_in), n, ctypes.byref(v), ctypes.byref(ms)))
                best = min(best, ms.value)
            extra[f"{name}_GBps"] = n / best / 1e6
            extra[f"{name}_frac_of_hbm_peak"] = n / best / 1e6 / peak
        # BASELINE.json configs[2]: level 1 on 1 GiB FASTQ-like
        fq = synthdata.make("fastq", 256 * MIB, 1)
        for r in range(4):
            ctx.h2d(d_in + r * fq.size, fq.ctypes.data_as(ctypes.c_void_p), fq.size)
        for lvl in (1, 6):
            best, o = 1e9, 0
            for _ in range(3):
                ms, o, _ = device_pass(level=lvl, nbytes=GIB, want_crc=False)
                best = min(best, ms)
            extra[f"L{lvl}_fastq_1GiB_GBps"] = GIB / best / 1e6
            extra[f"L{lvl}_fastq_ratio"] = o / GIB
        # CPU baseline beside it: the reference on all host cores, bounded sample of the same workload
        sample_n = min(n, a.cpu_sample_mib * MIB)
        sample = np.ascontiguousarray(np.tile(base, (sample_n + base.size - 1) // base.size)[:sample_n])
        dt, ref_out, kind, cores = cpu_reference_pass(sample, a.level, ncores)
        ctx.h2d(d_in, sample.ctypes.data_as(ctypes.c_void_p), sample_n)
        _, ours_out, _ = device_pass(nbytes=sample_n)
        cpu = {"value": sample_n / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": kind,
               "sample": f"first {sample_n // MIB} MiB of the workload, gzip_ng_threaded.open('wb', "
                         f"compresslevel={a.level}, threads={cores}, block_size=1 MiB) into a null sink",
               "ratio_reference": ref_out / sample_n, "ratio_ours": ours_out / sample_n,
               "size_ours_over_reference": ours_out / ref_out}
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": f"synthetic ({a.kind}: seeded {a.base_mib} MiB segment tiled to {a.gib:g} GiB per GPU)",
                "config": config, "ratio": ratio, "wall_ms_per_step": wall_ms / a.steps, "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roofline}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if extra:
            line["extra"] = extra
        emit_line(line)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
