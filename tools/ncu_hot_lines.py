#!/usr/bin/env python3
"""Hot source lines of one kernel from an ncu report captured with --import-source on:
    ncu -i REPORT --page source --csv --kernel-name regex:NAME --print-source cuda,sass > src.csv
    ncu_hot_lines.py src.csv [N]
Prints the N source lines with the most stall samples, with their share of the executed warp instructions."""
import csv
import sys


def num(x):
    try:
        return int(x.replace(",", ""))
    except ValueError:
        return 0


rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
idx = [k for k, r in enumerate(rows) if r and r[0] == "File Path"]
lines, tot = [], 0
for si, i in enumerate(idx):
    end = idx[si + 1] if si + 1 < len(idx) else len(rows)
    f = rows[i][1].split("/")[-1]
    for r in rows[i + 3:end]:
        if r and r[0].strip().isdigit():
            ie, samp = num(r[7]), num(r[4])
            lines.append((ie, samp, f, int(r[0]), r[1].strip()[:100]))
            tot += ie
stot = sum(l[1] for l in lines) or 1
print("warp instructions executed:", tot)
for ie, samp, f, ln, src in sorted(lines, key=lambda x: -x[1])[:top]:
    print("%5.1f%% inst %5.1f%% samples  %s:%d  %s" % (100 * ie / max(tot, 1), 100 * samp / stot, f, ln, src))
