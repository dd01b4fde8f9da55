#!/usr/bin/env python
"""SASS of a kernel's hot loop from one ncu capture, with executions per work unit and stall samples per instruction:
   python tools/ncu_sass_loop.py X.ncu-rep OPCODE:per_unit [min_ratio] > profiles/X_sass_row_loop.txt
(units as in tools/ncu_summary.py; instructions executed less than min_ratio times per unit are left out)."""
import csv
import subprocess
import sys

rep, unit = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.05
op, per = unit.rsplit(":", 1)
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
ix = {n: i for i, n in enumerate(rows[hi])}
data = []
for r in rows[hi + 1:]:
    try:
        data.append((int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]]), r[ix["Source"]].strip()))
    except (ValueError, IndexError):
        pass
def opcode(t):
    p = t.split()
    return p[1] if p[0].startswith("@") else p[0]
units = sum(n for n, _, t in data if opcode(t) == op) // int(per)
tot = sum(n for n, _, _ in data)
smp = sum(s for _, s, _ in data)
print("# %s: %d units (%s x %s per unit), %d warp instructions = %.1f per unit, %d stall samples" % (rep.split("/")[-1], units, op, per, tot, tot / units, smp))
print("# columns: SASS index, executions per unit, share of stall samples, instruction")
by_op = {}
for k, (n, s, t) in enumerate(data):
    if n >= thr * units:
        print("%5d %5.2f %5.2f%%  %s" % (k, n / units, 100.0 * s / smp, t))
    o = opcode(t).split(".")[0]
    by_op[o] = by_op.get(o, 0) + n
print("# warp instructions per unit by opcode family")
for o, n in sorted(by_op.items(), key=lambda kv: -kv[1])[:24]:
    print("#   %-12s %6.1f" % (o, n / units))
