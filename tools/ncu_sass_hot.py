#!/usr/bin/env python
"""Hot SASS instructions of an ncu source-page CSV (ncu -i X.ncu-rep --page source --csv --print-source sass):
   instruction counts, share of the kernel, stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ix = {n: i for i, n in enumerate(h)}
ie, isamp = ix["Instructions Executed"], ix["# Samples"]
data = []
for r in rows[hi + 1:]:
    try:
        data.append((int(r[ie]), int(r[isamp]), r))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data)
samp = sum(d[1] for d in data)
print("total warp instructions", tot, "samples", samp, "sass lines", len(data))
for k, (n, s, r) in enumerate(data):
    if n > tot * thr or s > samp * thr:
        print("%5d %12d %5.1f%% samp %5.1f%%  %s" % (k, n, 100.0 * n / tot, 100.0 * s / max(samp, 1), r[ix["Source"]].strip()[:100]))
