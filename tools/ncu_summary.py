#!/usr/bin/env python
"""Text summary of one `ncu --set full --import-source on` capture for profiles/: key counters (raw page), pipe
utilisation, stall reasons, share of samples / instructions per source file, the hottest source lines, and -- given the
number of work units of the launch (rows of a fill launch, nodes of a sort launch ...) -- warp instructions per unit.

    python tools/ncu_summary.py gpurun_out/X.ncu-rep [OPCODE:per_unit] > profiles/X_ncu_summary.txt

`OPCODE:per_unit` names a SASS opcode that is executed exactly `per_unit` times per unit and nowhere else in the kernel
(fill: VIMNMX3.S16x2:4 -- four per row); the sum of its execution counts / per_unit is the number of units."""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
unit = sys.argv[2] if len(sys.argv) > 2 else None

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__average_warp_latency_per_inst_issued.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg",
]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
if len(rows) >= 3:
    names, units, vals = rows[0], rows[1], rows[-1]
    d = {n: (u, v) for n, u, v in zip(names, units, vals)}
    print("kernel", d.get("Kernel Name", ("", "?"))[1])
    for k in KEYS:
        if k in d:
            print("%-70s %-14s %s" % (k, d[k][0], d[k][1]))
    for n in names:
        if n.startswith("sm__inst_executed_pipe_") and n.endswith(".sum") and n in d:
            print("%-70s %-14s %s" % (n, d[n][0], d[n][1]))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(src.splitlines()))
cur = None
hdr = None
lines = []
stalls = collections.Counter()
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 3 and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        dd = dict(zip(hdr, r))
        try:
            smp, ins = int(dd["# Samples"]), int(dd["Instructions Executed"])
        except (ValueError, KeyError):
            continue
        st = {}
        for k in hdr:
            if k.startswith("stall_") and "Not Issued" not in k:
                try:
                    st[k[6:]] = int(dd[k])
                except ValueError:
                    pass
        stalls.update(st)
        lines.append((cur, int(r[0]), r[1].strip(), smp, ins, st))
tot_s = sum(l[3] for l in lines) or 1
tot_i = sum(l[4] for l in lines) or 1
print("\ntotal samples %d, warp instructions %d" % (tot_s, tot_i))
n_units = None
if unit:
    op, per = unit.rsplit(":", 1)
    sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, text=True).stdout
    srows = list(csv.reader(sass.splitlines()))
    hi = next(i for i, r in enumerate(srows) if r and r[0] == "Address")
    ie, isrc = srows[hi].index("Instructions Executed"), srows[hi].index("Source")
    cnt = 0
    tot_sass = 0
    for r in srows[hi + 1:]:
        try:
            n = int(r[ie])
        except (ValueError, IndexError):
            continue
        tot_sass += n
        if r[isrc].strip().split()[0] == op or (r[isrc].strip().startswith("@") and r[isrc].split()[1] == op):
            cnt += n
    n_units = cnt // int(per)
    if n_units:
        print("units (%s executed %d times, %s per unit): %d  ->  %.1f warp instructions per unit (%d in all)" % (op, cnt, per, n_units, tot_sass / n_units, tot_sass))
print("\nstall reasons (share of samples)")
for k, v in stalls.most_common(10):
    print("  %-28s %5.1f%%" % (k, 100.0 * v / tot_s))
print("\nper source file")
per = collections.defaultdict(lambda: [0, 0])
for l in lines:
    per[l[0]][0] += l[3]
    per[l[0]][1] += l[4]
for f, (s, i) in sorted(per.items(), key=lambda kv: -kv[1][0]):
    if s or i:
        print("  %-34s %5.1f%% samples %5.1f%% instructions" % (f, 100.0 * s / tot_s, 100.0 * i / tot_i))
print("\nhottest source lines (by samples)")
for l in sorted(lines, key=lambda l: -l[3])[:40]:
    top = ",".join("%s:%d%%" % (k, round(100.0 * v / max(l[3], 1))) for k, v in sorted(l[5].items(), key=lambda kv: -kv[1])[:2])
    per_unit = (" %5.2f/unit" % (l[4] / n_units)) if n_units else ""
    print("  %-16s %4d %5.1f%% smp %5.1f%% inst%s  %-34s %s" % (l[0][:16], l[1], 100.0 * l[3] / tot_s, 100.0 * l[4] / tot_i, per_unit, top, l[2][:90]))
