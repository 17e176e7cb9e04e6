#!/usr/bin/env python3
"""Dev tool: top stall lines of a kernel from an .ncu-rep (SASS view). usage: ncu_src.py rep kernel_regex [N]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
N = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
si, ss, ie = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
body = []
for r in rows[h + 1:]:
    if not r or r[0] in ("Address", "Kernel Name"):
        break
    if len(r) == len(hdr):
        body.append(r)
tot = sum(int(r[ss] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
top = sorted(range(len(body)), key=lambda i: -int(body[i][ss] or 0))[:N]
for i in sorted(top):
    r = body[i]
    st = sorted(((int(r[c] or 0), hdr[c][6:]) for c in stall_cols), reverse=True)[:2]
    print(f"{i:4d} {int(r[ss] or 0):6d} {100*int(r[ss] or 0)/max(tot,1):5.1f}% exec={r[ie]:>9s} {r[si][:70]:70s} {st}")
