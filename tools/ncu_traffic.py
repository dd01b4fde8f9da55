"""profiles/ncu_traffic.json from an ncu launch list (the `--metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum[,smsp__inst_executed.sum] --csv` pass of bench.py): per kernel, the launches of ONE pass over the
batch (from one poa_init_kernel to the next; the first complete pass of the capture), their DRAM bytes, their duration
under ncu (cold-cache, serialised: shares, not absolutes) and the bytes per window and step.

    python tools/ncu_traffic.py gpurun_out/X_launches.csv 18944 > profiles/ncu_traffic.json
"""
import collections
import csv
import json
import sys


def main():
    path, n_windows = sys.argv[1], int(sys.argv[2])
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hi]
    ix = {n: i for i, n in enumerate(h)}
    launches = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) < len(h):
            continue
        lid = int(r[ix["ID"]])
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("poa_b200::", "").split("<")[0]
        v = float(r[ix["Metric Value"]].replace(",", ""))
        unit = r[ix["Metric Unit"]]
        m = r[ix["Metric Name"]]
        if m == "gpu__time_duration.sum":
            v = v / 1e6 if unit == "ns" else v / 1e3 if unit == "us" else v if unit == "ms" else v * 1e3
        elif m.startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        launches.setdefault(lid, {"name": name})[m] = v
    ids = sorted(launches)
    inits = [i for i in ids if launches[i]["name"] == "poa_init_kernel"]
    if len(inits) < 2:
        raise SystemExit("the capture does not hold one complete pass (two poa_init_kernel launches)")
    lo, hi_ = inits[0], inits[1]
    out = collections.OrderedDict()
    total = 0.0
    for i in ids:
        if not (lo <= i < hi_):
            continue
        L = launches[i]
        k = out.setdefault(L["name"], {"launches": 0, "dram_bytes_read": 0.0, "dram_bytes_write": 0.0, "kernel_ms": 0.0,
                                       "warp_instructions": 0.0})
        k["launches"] += 1
        k["dram_bytes_read"] += L.get("dram__bytes_read.sum", 0.0)
        k["dram_bytes_write"] += L.get("dram__bytes_write.sum", 0.0)
        k["kernel_ms"] += L.get("gpu__time_duration.sum", 0.0)
        k["warp_instructions"] += L.get("smsp__inst_executed.sum", 0.0)
    for k in out.values():
        b = k["dram_bytes_read"] + k["dram_bytes_write"]
        k["dram_bytes_per_window_step"] = b / n_windows
        k["dram_gb_per_s_under_ncu"] = b / (k["kernel_ms"] * 1e-3) / 1e9 if k["kernel_ms"] > 0 else None
        total += b
    out["_total"] = {"windows": n_windows, "dram_bytes_per_window_step": total / n_windows,
                     "kernel_ms": sum(k["kernel_ms"] for k in out.values()),
                     "source": path.split("/")[-1] + ": one pass of bench.py's batch under ncu "
                               "(--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none)"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
