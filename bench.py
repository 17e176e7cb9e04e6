#!/usr/bin/env python3
"""bench.py — headline benchmark: gzip/DEFLATE compress throughput (input GB/s).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle/_ref)

Workload (BASELINE.json configs[3]): level-6 compression of ONE stream of 4·N GiB of mixed-entropy data, 4 GiB per
GPU, in 128 KiB device blocks primed with the previous 32 KiB; one "step" = one pass of the whole pipeline over the
stream.  The stream shards by contiguous range (zlib_ng_b200.sharding): rank r holds shard r plus the 32 KiB of shard
r-1 before it, no collective touches the data path ("weak" scaling); torch.distributed (NCCL) carries the barrier, the
max-over-ranks of the device times and the gather of the per-shard (size, crc32, length) triples that rank 0 folds
with crc32_combine — the reference writer's fold (gzip_ng_threaded.py:394).

value   device-resident: every shard already in HBM, output left in HBM, CUDA events, max over ranks
e2e     the same stream from pinned HOST memory to one DEFLATE stream in pinned host memory through the C ABI's
        multi-device call (b2d_compress_and_crc_multi: one process, N GPUs, H2D + kernels + CRC + D2H inside the timed
        region), driven by rank 0; verified in the run (`verify_ok`)
"""
import argparse
import ctypes
import json
import os
import subprocess
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
BLOCK = 128 * 1024
WINDOW = 32768
METRIC = "gzip_compress_input_GBps_L6"
STAGES = ["k_chains", "k_match", "k_parse", "k_plan", "k_scan", "k_emit"]
# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per INPUT byte of a launch, from the `ncu --set full`
# capture of a 256 MiB level-6 batch of this workload (profiles/README.md names the capture the figures come from)
NCU_DRAM_BYTES_PER_INPUT_BYTE = {"k_chains": 3.58, "k_match": 7.33, "k_parse": 8.97, "k_plan": 0.063,
                                 "k_scan": 0.004, "k_emit": 2.11}
NCU_SOURCE = "profiles/ncu_r2_summary.txt"


# ---------------------------------------------------------------------------------------------- workload
def shard_view(base: np.ndarray, r: int):
    """Shard r of the logical stream = the seeded base segment rotated by r * 8 MiB + r * 4099 bytes, tiled.  Every
    shard holds the same kinds of data (same work per byte) but different bytes, so the per-shard CRCs differ and a
    fold in the wrong order or with a wrong length is caught."""
    rot = (r * (8 * MIB + 4099)) % base.size
    return rot


def shard_bytes(base: np.ndarray, r: int, n: int) -> np.ndarray:
    """shard r as a new array: period = the rotated base"""
    rot = shard_view(base, r)
    period = np.concatenate([base[rot:], base[:rot]]) if rot else base
    reps = (n + period.size - 1) // period.size
    return np.tile(period, reps)[:n] if reps > 1 else period[:n].copy()


def shard_tail(base: np.ndarray, r: int, n: int, k: int = WINDOW) -> np.ndarray:
    """the last k bytes of shard r without building it"""
    rot = shard_view(base, r)
    idx = (np.arange(n - k, n, dtype=np.int64) + rot) % base.size
    return base[idx]


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons of one GPU every 20 ms while the timed region runs."""

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


def dist_setup():
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
        global _CPU_GROUP
        _CPU_GROUP = dist_mod.new_group(backend="gloo")
    return rank, world, local, dist


_CPU_GROUP = None


def cpu_barrier(dist):
    """A barrier that leaves the GPUs alone.  While rank 0 drives all N GPUs from one process (the end-to-end leg, the
    streaming legs) the other ranks must not sit in an NCCL barrier: its kernel spins on their GPU, and kernels of two
    processes on one GPU are time-sliced, which halves what rank 0 gets there (measured: 36.7 instead of 77 GB/s at N=2)."""
    if dist is not None:
        dist.barrier(group=_CPU_GROUP)


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


def gather_triples(dist, local, world, triple):
    """every rank's (compressed bytes, crc32, input bytes) on every rank, in rank order"""
    if dist is None:
        return [triple]
    import torch
    mine = torch.tensor([int(x) for x in triple], dtype=torch.int64, device=f"cuda:{local}")
    out = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return [tuple(int(v) for v in t.tolist()) for t in out]


def reference_module():
    pkg = os.path.join(ROOT, "oracle", "_ref", "pkg")
    if os.path.isdir(os.path.join(pkg, "zlib_ng")):
        if pkg not in sys.path:
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


def cpu_reference_pass(sample: np.ndarray, level: int, threads: int, block_size: int = 1 << 20):
    """The reference's own hot path on the host cores: gzip_ng_threaded.open('wb', threads=all).
    Falls back to the repo's CPU twin (a port of OUR pipeline, single thread) if oracle/_ref is absent."""
    ref = reference_module()
    mv = memoryview(sample)
    if ref is not None:
        _, gzt = ref
        sink = NullSink()
        t0 = time.perf_counter()
        with gzt.open(sink, "wb", compresslevel=level, threads=threads, block_size=block_size) as f:
            for i in range(0, sample.size, 8 * MIB):
                f.write(mv[i:i + 8 * MIB])
        dt = time.perf_counter() - t0
        return dt, sink.n, "reference", threads
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import twin_deflate
    t0 = time.perf_counter()
    out = twin_deflate(sample.tobytes(), level)
    return time.perf_counter() - t0, len(out), "port", 1


def run_tool(args, timeout):
    """one JSON line from a tool run in a FRESH process (cold-start numbers need one)"""
    try:
        p = subprocess.run([sys.executable] + args, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
        lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
        if p.returncode == 0 and lines:
            return json.loads(lines[-1])
        return {"error": f"rc={p.returncode}: {p.stderr[-300:]}"}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"}


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
    ap.add_argument("--no-stream", action="store_true", help="skip the configs[0] / configs[4] legs (fresh subprocesses)")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    n = int(a.gib * GIB) // BLOCK * BLOCK
    ncores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    # identical in both arms (the driver compares the two dicts)
    config = {"workload": f"level-{a.level} DEFLATE of one {a.gib * a.gpus:g} GiB {a.kind}-entropy synthetic stream, "
                          f"{a.gib:g} GiB per GPU, 128 KiB blocks + 32 KiB dictionary (BASELINE.json configs[3]); "
                          f"a seeded {a.base_mib} MiB segment, rotated per shard and tiled (the tile period is far beyond "
                          f"DEFLATE's 32 KiB window, so tiling changes neither the work per byte nor the ratio)",
              "bytes_per_gpu": n, "level": a.level, "block": BLOCK, "batch_mib": a.batch_mib,
              "l2": "inputs (GiB) far larger than the 126 MB L2; no flush needed",
              "parallelism": f"{a.gpus} contiguous shard(s) of one stream, no collective on the data path"}

    if a.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        if rank != 0:
            return 0
        sample_n = min(n, a.cpu_sample_mib * MIB)
        base = synthdata.make(a.kind, min(n, a.base_mib * MIB), 2)
        sample = np.ascontiguousarray(shard_bytes(base, 0, sample_n))
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
                                 "sample": f"first {sample_n // MIB} MiB of the stream per step through "
                                           f"gzip_ng_threaded.open('wb', compresslevel={a.level}, threads={cores}), "
                                           f"ratio {out_n / sample_n:.4f}"},
                "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit_line(line)
        return 0

    # ------------------------------------------------------------------ this repo's arm
    claim_stdout()
    rank, world, local, dist = dist_setup()
    from zlib_ng_b200 import _native as nat
    from zlib_ng_b200 import sharding
    # the logical stream and this rank's shard of it: [32 KiB of the previous shard | shard]
    base = synthdata.make(a.kind, min(n, a.base_mib * MIB), 2)
    plan = sharding.plan_shards(world * n, world)
    assert plan[rank] == (rank * n, (rank + 1) * n)
    zdict = shard_tail(base, rank - 1, n) if rank else np.empty(0, dtype=np.uint8)
    dn = int(zdict.size)
    L = nat.lib()
    ctx = nat.Context(local, a.batch_mib * MIB)
    d_in = ctx.device_alloc(dn + n + 64)
    cap = (int(L.b2d_deflate_bound(n)) + 3) & ~3
    d_out = ctx.device_alloc(cap)
    shard = shard_bytes(base, rank, n)
    if dn:
        ctx.h2d(d_in, zdict.ctypes.data_as(ctypes.c_void_p), dn)
    for off in range(0, n, GIB):
        m = min(GIB, n - off)
        ctx.h2d(d_in + dn + off, ctypes.c_void_p(shard.ctypes.data + off), m)

    def device_pass(level=a.level, nbytes=n, want_crc=True, dict_n=dn):
        out_n, crc, ms = ctypes.c_uint64(0), ctypes.c_uint32(0), ctypes.c_float(0)
        nat.check(L.b2d_deflate_device(ctx.handle, level, ctypes.c_void_p(d_in + (dn - dict_n)), dict_n, nbytes,
                                       ctypes.c_void_p(d_out), cap, ctypes.byref(out_n),
                                       ctypes.byref(crc) if want_crc else None, ctypes.byref(ms)))
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

    # one stream: rank 0 folds the shards' (size, crc, length) in order, as the reference's writer thread does, and
    # checks the fold against the reference's crc32 of every shard combined by the REFERENCE's crc32_combine
    ref = reference_module()
    import zlib
    crc_mod = ref[0] if ref is not None else zlib
    host_crc = 0
    smv = memoryview(shard)
    for off in range(0, n, GIB):
        host_crc = crc_mod.crc32(smv[off:off + GIB], host_crc)
    triples = gather_triples(dist, local, world, (out_n, crc, n))
    host_triples = gather_triples(dist, local, world, (0, host_crc, n))
    stream_out, stream_crc, stream_n = 0, 0, 0
    want_crc = 0
    for (o, c, ln), (_, hc, _) in zip(triples, host_triples):
        stream_crc = int(L.b2d_crc32_combine(stream_crc, c, ln))
        want_crc = crc_mod.crc32_combine(want_crc, hc, ln) if ref is not None else int(L.b2d_crc32_combine(want_crc, hc, ln))
        stream_out += o
        stream_n += ln
    shards_ok = stream_crc == want_crc and all(c == hc for (_, c, _), (_, hc, _) in zip(triples, host_triples))
    ratio = stream_out / stream_n

    # roofline of the dominant kernel (largest share of the step), from the live per-kernel events
    shares = {STAGES[i]: stage_ms[i] for i in range(6)}
    top = max(range(6), key=lambda i: stage_ms[i])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes_per_launch = (1.0 + out_n / n) * n * a.steps / max(int(stage_n[top]), 1)
    achieved = alg_bytes_per_launch / (stage_ms[top] / max(int(stage_n[top]), 1) * 1e-3) / 1e9
    achieved_alone = alg_bytes_per_launch / (alone_ms[top] / max(int(stage_n[top]) // a.steps, 1) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": STAGES[top], "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "achieved_alone": achieved_alone, "frac_alone": achieved_alone / peak,
                "durations": "achieved: CUDA events around the kernel inside the timed region, where two batches overlap on two "
                             "streams (a stage's time includes what it shares the SMs with); achieved_alone: the same kernel "
                             "with one batch in flight (b2d_profile mode 2, one pass after the timed region)",
                "traffic": NCU_DRAM_BYTES_PER_INPUT_BYTE[STAGES[top]] * n * a.steps / max(int(stage_n[top]), 1),
                "traffic_source": f"ncu --set full capture of one launch ({NCU_SOURCE}), DRAM bytes per "
                                  "input byte x input bytes per launch",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
                "algorithmic_bytes_per_launch": alg_bytes_per_launch,
                "kernel_ms_per_step": {k: v / a.steps for k, v in shares.items()},
                "kernel_ms_per_step_alone": {STAGES[i]: alone_ms[i] for i in range(6)},
                "note": "DEFLATE is bound by shared-memory / integer work, not HBM; see DESIGN.md §4"}

    # ------------------------------------------------------------------ end to end: ONE stream, host buffers, all N GPUs
    e2e, verify = None, {}
    ctx.device_free(d_in)
    ctx.device_free(d_out)
    d_in = d_out = None
    barrier(dist, local)
    ctx.sync()
    if rank == 0:
        try:
            total = world * n
            cap_all = int(L.b2d_deflate_bound(total)) + 64
            h_in, h_out = nat.pinned_alloc(total), nat.pinned_alloc(cap_all)
            hv = np.frombuffer((ctypes.c_uint8 * total).from_address(h_in), dtype=np.uint8)
            for r in range(world):
                hv[r * n:(r + 1) * n] = shard if r == rank else shard_bytes(base, r, n)
            # several GPUs: 64 MiB batches — four lanes of them fill a device like two of 256 MiB do, with a quarter of the
            # pipeline's fill and drain time (profiles/r2_i_multi_e2e.log: 100.7 against 95.9 GB/s on 8 GPUs; one GPU:
            # 39.1 against 40.4)
            m = nat.MultiContext(list(range(world)), (a.batch_mib if world == 1 else min(a.batch_mib, 64)) * MIB)

            def host_pass():
                return m.compress_and_crc(a.level, h_in, total, None, 0, h_out, cap_all, want_sizes=True)
            host_pass()
            l0 = m.launch_count()
            t0 = time.perf_counter()
            for _ in range(a.steps):
                o_n, c_e2e, sizes = host_pass()
            t_e2e = time.perf_counter() - t0
            e2e_launches = m.launch_count() - l0
            e2e = {"value": total * a.steps / t_e2e / 1e9, "unit": "GB/s", "h2d_bytes_per_step": total,
                   "d2h_bytes_per_step": o_n, "ms_per_step": 1e3 * t_e2e / a.steps, "gpu_launches": int(e2e_launches),
                   "api": f"b2d_compress_and_crc_multi(one {total / GIB:g} GiB stream in pinned host memory -> one DEFLATE "
                          f"stream in pinned host memory, {world} GPU(s), one process), the call behind gzip_ng_threaded"}
            # ---- verification of this very output (outside the timed region)
            ov = np.frombuffer((ctypes.c_uint8 * o_n).from_address(h_out), dtype=np.uint8)
            verify["stream_crc_equals_reference_crc32"] = c_e2e == want_crc
            verify["stream_size_equals_sum_of_shards"] = o_n == stream_out and sum(sizes) == o_n
            # the N-GPU stream is byte-identical to what ONE context produces for the same bytes: a 1 GiB window that
            # starts in shard 0 and (N > 1) ends in shard 1, i.e. crosses a shard and many batch boundaries
            win0 = max(0, n - GIB // 2) // BLOCK * BLOCK if world > 1 else 0
            win_n = min(GIB, total - win0)
            b0, nb = win0 // BLOCK, (win_n + BLOCK - 1) // BLOCK
            o0, o1 = sum(sizes[:b0]), sum(sizes[:b0 + nb])
            one = nat.Context(0, a.batch_mib * MIB)
            cap1 = int(L.b2d_deflate_bound(win_n)) + 64
            buf1 = ctypes.create_string_buffer(cap1)
            on1, c1 = ctypes.c_uint64(0), ctypes.c_uint32(0)
            dict0 = hv[win0 - WINDOW:win0] if win0 else None
            nat.check(L.b2d_compress_and_crc(one.handle, a.level, ctypes.c_void_p(h_in + win0), win_n,
                                             ctypes.c_void_p(h_in + win0 - WINDOW) if win0 else None, WINDOW if win0 else 0,
                                             buf1, cap1, ctypes.byref(on1), ctypes.byref(c1)))
            one.close()
            piece = ov[o0:o1].tobytes()
            verify["identical_to_one_gpu_on_1GiB_window_across_shards"] = piece == buf1.raw[:on1.value]
            # ... and the REFERENCE's decoder gives the input back (first 256 MiB of that window: 0.4 GB/s on one core)
            dec_n = min(256 * MIB, win_n)
            db = dec_n // BLOCK
            dpiece = ov[o0:o0 + sum(sizes[b0:b0 + db])].tobytes() + b"\x03\x00"
            dmod = ref[0] if ref is not None else zlib
            d = dmod.decompressobj(-15, zdict=dict0.tobytes()) if dict0 is not None else dmod.decompressobj(-15)
            verify["reference_decoder_round_trip_256MiB"] = d.decompress(dpiece) == hv[win0:win0 + dec_n].tobytes()
            verify["decoder"] = "reference zlib_ng.decompressobj" if ref is not None else "stdlib zlib"
            m.close()
            nat.pinned_free(h_in)
            nat.pinned_free(h_out)
        except Exception as e:  # report, never fake
            e2e = {"value": None, "unit": "GB/s", "error": f"{type(e).__name__}: {e}"}
    cpu_barrier(dist)             # ranks > 0 sleep here (no GPU work) while rank 0 drives every GPU
    verify["shard_crcs_fold_to_reference_crc32"] = bool(shards_ok)
    verify_ok = bool(e2e and e2e.get("value")) and all(v for k, v in verify.items() if isinstance(v, bool))

    # ------------------------------------------------------------------ extras + CPU baseline (rank 0)
    extra, cpu = {}, None
    if rank == 0 and not a.no_extras and not a.no_stream:
        devs = ",".join(str(i) for i in range(world))
        tool = os.path.join(ROOT, "tools", "stream_bench.py")
        # BASELINE.json configs[4] (16 GiB streaming write through the Python API, all N GPUs, cold and warm, the reference
        # beside it at threads = host cores) and configs[0] (256 MiB FASTQ-like, level 1, threads=4), each in a fresh process
        extra["configs4_stream_16GiB"] = run_tool([tool, "--config", "4", "--devices", devs], 600)
        extra["configs0_fastq_256MiB_L1_t4"] = run_tool([tool, "--config", "0", "--devices", "0"], 300)
    if rank == 0 and world == 1 and not a.no_extras:
        d_in = ctx.device_alloc(n + 64)
        d_out = ctx.device_alloc(cap)
        dn = 0
        for off in range(0, n, GIB):
            ctx.h2d(d_in + off, ctypes.c_void_p(shard.ctypes.data + off), min(GIB, n - off))
        # correctness of this very run: round-trip of the first 64 MiB + CRC, through the reference's decoder
        head_n = 64 * MIB
        _, o1, c1 = device_pass(nbytes=head_n, dict_n=0)
        host = np.empty(o1, dtype=np.uint8)
        ctx.d2h(host.ctypes.data_as(ctypes.c_void_p), d_out, o1)
        want = shard[:head_n].tobytes()
        dmod = ref[0] if ref is not None else zlib
        extra["roundtrip_64MiB_ok"] = dmod.decompressobj(-15).decompress(host.tobytes() + b"\x03\x00") == want
        extra["roundtrip_decoder"] = "reference zlib_ng.decompressobj" if ref is not None else "stdlib zlib"
        extra["crc_ok"] = c1 == crc_mod.crc32(want)
        # the other side of the wire: block-parallel inflate of this run's own output,
        # device-resident, checked with the CRC kernel; beside it the reference's inflate (one thread, as its
        # reader runs it) on the first 256 MiB worth of the same stream
        try:
            _, o_full, c_full = device_pass(dict_n=0)
            cnt = ctypes.c_uint64(0)
            nat.check(L.b2d_block_index(ctx.handle, None, 0, ctypes.byref(cnt)))
            offs = (ctypes.c_uint64 * cnt.value)()
            nat.check(L.b2d_block_index(ctx.handle, offs, cnt.value, ctypes.byref(cnt)))
            nbk = cnt.value - 1
            bits = (ctypes.c_uint32 * (4 * nbk))()
            nat.check(L.b2d_subblock_bits(ctx.handle, bits, 4 * nbk, ctypes.byref(cnt)))
            d_back = ctx.device_alloc(n + 64)
            for tag, sub in (("inflate", None), ("inflate_sub_blocks", bits)):      # one warp per block / per 32 KiB sub-block
                best = 1e9
                for _ in range(3):
                    ims = ctypes.c_float(0)
                    nat.check(L.b2d_inflate_device_sub(ctx.handle, ctypes.c_void_p(d_out), o_full, offs, sub, nbk, None, 0,
                                                       ctypes.c_void_p(d_back), n, ctypes.byref(ims)))
                    best = min(best, ims.value)
                v = ctypes.c_uint32(0)
                nat.check(L.b2d_crc32_device(ctx.handle, 0, ctypes.c_void_p(d_back), n, ctypes.byref(v), None))
                extra[f"{tag}_GBps"] = n / best / 1e6
                extra[f"{tag}_crc_ok"] = v.value == c_full
            ctx.device_free(d_back)
            cnt.value = nbk + 1
            nblk = min(2048, cnt.value - 1)
            comp = np.empty(offs[nblk], dtype=np.uint8)
            ctx.d2h(comp.ctypes.data_as(ctypes.c_void_p), d_out, int(offs[nblk]))
            t0 = time.perf_counter()
            back_n = len(dmod.decompressobj(-15).decompress(comp.tobytes()))
            extra["inflate_reference_1thread_GBps"] = back_n / (time.perf_counter() - t0) / 1e9
            extra["inflate_reference_kind"] = "reference zlib_ng.decompressobj" if ref is not None else "stdlib zlib"
        except Exception as e:  # report, never fake
            extra["inflate_error"] = f"{type(e).__name__}: {e}"
        ms1 = min(device_pass(level=1, dict_n=0)[0] for _ in range(3))
        extra["L1_mixed_GBps"] = n / ms1 / 1e6
        for name, fn, start in (("crc32", L.b2d_crc32_device, 0), ("adler32", L.b2d_adler32_device, 1)):
            best = 1e9
            for _ in range(4):
                v, ms = ctypes.c_uint32(0), ctypes.c_float(0)
                nat.check(fn(ctx.handle, start, ctypes.c_void_p(d_in), n, ctypes.byref(v), ctypes.byref(ms)))
                best = min(best, ms.value)
            extra[f"{name}_GBps"] = n / best / 1e6
            extra[f"{name}_frac_of_hbm_peak"] = n / best / 1e6 / peak
        # CPU baseline beside it: the reference on all host cores, bounded sample of the same workload, at the
        # reference's default block_size (1 MiB) and at our device-block size (128 KiB: "same block_size")
        sample_n = min(n, a.cpu_sample_mib * MIB)
        sample = np.ascontiguousarray(shard[:sample_n])
        dt, ref_out, kind, cores = cpu_reference_pass(sample, a.level, ncores)
        _, ref_out_128k, _, _ = cpu_reference_pass(sample[:min(sample_n, 256 * MIB)], a.level, ncores, 128 * 1024)
        _, ours_out, _ = device_pass(nbytes=sample_n, dict_n=0)
        _, ours_256, _ = device_pass(nbytes=min(sample_n, 256 * MIB), dict_n=0)
        cpu = {"value": sample_n / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": kind,
               "sample": f"first {sample_n // MIB} MiB of the workload, gzip_ng_threaded.open('wb', "
                         f"compresslevel={a.level}, threads={cores}, block_size=1 MiB) into a null sink",
               "ratio_reference": ref_out / sample_n, "ratio_ours": ours_out / sample_n,
               "size_ours_over_reference": ours_out / ref_out,
               "size_ours_over_reference_at_block_size_128KiB": ours_256 / ref_out_128k}
        # BASELINE.json configs[2]: level 1 on 1 GiB FASTQ-like; sizes next to the reference's at the same levels,
        # measured here (256 MiB of the same data, block_size 1 MiB and 128 KiB)
        fq = synthdata.make("fastq", 256 * MIB, 1)
        for r in range(4):
            ctx.h2d(d_in + r * fq.size, fq.ctypes.data_as(ctypes.c_void_p), fq.size)
        for lvl in (1, 6):
            best, o = 1e9, 0
            for _ in range(3):
                ms, o, _ = device_pass(level=lvl, nbytes=GIB, want_crc=False, dict_n=0)
                best = min(best, ms)
            extra[f"L{lvl}_fastq_1GiB_GBps"] = GIB / best / 1e6
            extra[f"L{lvl}_fastq_ratio"] = o / GIB
            if ref is not None:
                _, o256, _ = device_pass(level=lvl, nbytes=fq.size, want_crc=False, dict_n=0)
                for bs, tag in ((1 << 20, "1MiB"), (128 * 1024, "128KiB")):
                    rdt, rout, _, _ = cpu_reference_pass(fq, lvl, ncores, bs)
                    extra[f"L{lvl}_fastq_ratio_reference_block_{tag}"] = rout / fq.size
                    extra[f"L{lvl}_fastq_size_ours_over_reference_block_{tag}"] = o256 / rout
                    if bs == 1 << 20:
                        extra[f"L{lvl}_fastq_reference_GBps_{ncores}_cores"] = fq.size / rdt / 1e9
        # the user-facing call, warm, in this process: pageable bytes in 64 MiB writes, and one write from a pinned_buffer
        try:
            import io
            from zlib_ng_b200 import gzip_ng_threaded as gzt

            class _Null(io.RawIOBase):
                n = 0

                def writable(self):
                    return True

                def write(self, b):
                    self.n += len(b)
                    return len(b)
            os.environ["B2D_DEVICES"] = "0"
            smv = memoryview(shard)
            best = 0.0
            for _ in range(3):
                sink = _Null()
                t0 = time.perf_counter()
                with gzt.open(sink, "wb", compresslevel=a.level, threads=ncores) as f:
                    for off in range(0, n, 64 * MIB):
                        f.write(smv[off:off + 64 * MIB])
                best = max(best, n / (time.perf_counter() - t0) / 1e9)
            extra["gzip_ng_threaded_open_wb_GBps"] = best
            extra["gzip_ng_threaded_ratio"] = sink.n / n
            with gzt.pinned_buffer(n) as pb:
                np.frombuffer(pb.view, dtype=np.uint8)[:] = shard
                best = 0.0
                for _ in range(3):
                    sink = _Null()
                    t0 = time.perf_counter()
                    with gzt.open(sink, "wb", compresslevel=a.level, threads=ncores) as f:
                        f.write(pb.view)
                    best = max(best, n / (time.perf_counter() - t0) / 1e9)
            extra["gzip_ng_threaded_open_wb_pinned_buffer_GBps"] = best
            # ... and the way back: the file (tmpfs) through open(..., "rb") on the GPU, readinto a reused 64 MiB buffer;
            # the bytes are checked (CRC-32 by the reader itself, against the trailer; a wrong length raises too)
            shm = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
            path = os.path.join(shm, f"b2d_bench_{os.getpid()}.gz")
            try:
                with gzt.open(path, "wb", compresslevel=a.level, threads=ncores) as f:
                    f.write(smv)
                box = bytearray(64 * MIB)
                best, on_gpu = 0.0, False
                for _ in range(3):
                    got = 0
                    t0 = time.perf_counter()
                    with gzt.open(path, "rb") as f:
                        on_gpu = f.raw._gpu is not None
                        while True:
                            k = f.readinto(box)
                            if not k:
                                break
                            got += k
                    if got != n:
                        raise RuntimeError(f"reader returned {got} of {n} bytes")
                    best = max(best, n / (time.perf_counter() - t0) / 1e9)
                extra["gzip_ng_threaded_open_rb_GBps"] = best
                extra["gzip_ng_threaded_open_rb_on_gpu"] = on_gpu
            finally:
                if os.path.exists(path):
                    os.unlink(path)
        except Exception as e:  # report, never fake
            extra["gzip_ng_threaded_error"] = f"{type(e).__name__}: {e}"
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": f"synthetic ({a.kind}: seeded {a.base_mib} MiB segment, rotated per shard, tiled to "
                                       f"{a.gib:g} GiB per GPU)",
                "config": config, "ratio": ratio, "wall_ms_per_step": wall_ms / a.steps, "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roofline, "verify_ok": verify_ok, "verify": verify,
                "stream": {"bytes": stream_n, "deflate_bytes": stream_out, "crc32": stream_crc, "shards": world}}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if extra:
            line["extra"] = extra
        emit_line(line)
    cpu_barrier(dist)             # the other ranks leave together with rank 0 (its extras run last)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
