#!/usr/bin/env python3
"""Turns the round's ncu captures (tools/run_ncu_r2.sh) into the text committed under profiles/:
    python tools/ncu_summary.py gpurun_out/launches_r2.csv gpurun_out/prof_r2_raw.csv > profiles/ncu_r2_summary.txt"""
import csv
import re
import sys
from collections import OrderedDict

launches, raw = sys.argv[1], sys.argv[2]
METRICS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
           "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed"]


def short(name):
    return re.sub(r"\(.*", "", name.replace("void ", "")).strip()


tot = OrderedDict()
with open(launches) as f:
    rows = [r for r in csv.reader(f) if len(r) > 14 and r[0].isdigit()]
for r in rows:
    k = short(r[4])
    n, ns = tot.get(k, (0, 0.0))
    tot[k] = (n + 1, ns + float(r[14].replace(",", "")))
allns = sum(v[1] for v in tot.values())
print("1. Launch list: ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --steps 2 --warmup 3 --no-extras")
print("   (cold-cache, serialised launches: the SHARES are what counts, the bench's own CUDA events give the times)")
print(f"   {'kernel':<22}{'launches':>9}{'total ms':>12}{'share':>8}")
for k, (n, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"   {k:<22}{n:>9}{ns / 1e6:>12.2f}{ns / allns:>8.3f}")
print()
print("2. ncu --set full --clock-control none --import-source on, one launch per kernel on one 256 MiB level-6 batch of the bench")
print("   workload (tools/ncu_target.py); the inflate kernels on the stream that batch produced")
with open(raw) as f:
    rr = list(csv.reader(f))
hdr, units = rr[0], rr[1]
ki = hdr.index("Kernel Name")
seen = set()
for r in rr[2:]:
    k = short(r[ki])
    if k in seen:
        continue
    seen.add(k)
    print(f"== {k}")
    for m in METRICS:
        cols = [i for i, h in enumerate(hdr) if h == m or h.endswith("." + m)]
        if cols:
            i = cols[0]
            print(f"   {m:<82}{r[i]:>16} {units[i]}")
