"""GPU box: the native engine (both device schedules) against the UNMODIFIED reference cudapoa build
(oracle/_ref/libcudapoa_ref.so, test infrastructure) on many DISTINCT windows -- statuses, consensus, coverage or MSA
rows compared window by window.  Reads nothing from /root/reference.

    python tools/parity_at_scale.py [out.json]

Cases: cfg 2 windows (plain and with PHRED-like weights), the same with MSA output, noisy reads, band 128, mixed sizes
(cfg 5b: int32 scores, rebased packed rows) and long windows (cfg 4).  Sizes are chosen so that the reference build
finishes each case in seconds to a minute."""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.windows import config_windows, generate_windows  # noqa: E402

# name: (window factory, engine parameters, memory, windows)
CASES = [
    ("cfg2_consensus", lambda n: config_windows("cfg2", n, start_index=100000),
     dict(max_seq=640, max_reads=30, band=256, banded=True, output="consensus"), 24 << 30, 12288),
    ("cfg2_weighted", lambda n: config_windows("cfg2", n, start_index=200000, with_weights=True),
     dict(max_seq=640, max_reads=30, band=256, banded=True, output="consensus"), 24 << 30, 4096),
    ("cfg3_msa", lambda n: config_windows("cfg3", n, start_index=300000),
     dict(max_seq=640, max_reads=30, band=256, banded=True, output="msa"), 24 << 30, 2048),
    ("noisy_25pct", lambda n: generate_windows(n, start_index=400000, template_length=600, n_reads=20, read_error=0.25,
                                               backbone_error=0.15),
     dict(max_seq=896, max_reads=20, band=256, banded=True, output="consensus"), 24 << 30, 2048),
    ("band128", lambda n: generate_windows(n, start_index=500000, template_length=500, n_reads=12),
     dict(max_seq=640, max_reads=12, band=128, banded=True, output="consensus"), 16 << 30, 2048),
    ("cfg5b_mixed", lambda n: config_windows("cfg5b", n, start_index=600000),
     dict(max_seq=2560, max_reads=60, band=256, banded=True, output="consensus"), 100 << 30, 2048),
    ("cfg4_long", lambda n: config_windows("cfg4", n, start_index=700000),
     dict(max_seq=6400, max_reads=100, band=256, banded=True, output="consensus"), 60 << 30, 48),
]


def digest(res, output):
    h = hashlib.sha256()
    for w in range(len(res["status"])):
        h.update(parity.window_digest(res, w, output).encode())
    return h.hexdigest()[:20]


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "parity_at_scale.json")
    only = set(sys.argv[2:])
    nat = _cabi.native()
    ref = parity.ref_lib()
    report = {"reference": ref.implementation().decode(), "native": nat.implementation().decode(), "cases": []}
    all_ok = True
    for name, make, eng, mem, n in CASES:
        if only and name not in only:
            continue
        ws = make(n)
        output = eng["output"]
        row = {"case": name, "windows": n, "reads": int(ws.n_reads), "bases": int(ws.n_bases), "engine": eng}
        t0 = time.time()
        r = parity.run_abi(ref, ws, max_mem=mem, **eng)
        row["reference_seconds"] = round(time.time() - t0, 2)
        row["statuses"] = {str(k): int(v) for k, v in zip(*np.unique(np.asarray(r["status"], dtype=np.int64), return_counts=True))}
        row["reference_digest"] = digest(r, output)
        for mode, label in ((1, "single_launch"), (2, "phase_split")):
            nat._dll.cudapoa_b200_debug_set_schedule(mode)
            t0 = time.time()
            g = parity.run_abi(nat, ws, max_mem=mem, **eng)
            bad = parity.compare(g, r, output)
            row[label] = {"seconds": round(time.time() - t0, 2), "mismatching_windows": len(bad), "first": bad[:8],
                          "exact_path_alignments": int(sum(s["exact_path"] for s in g["stats"])),
                          "alignments": int(sum(s["reads_aligned"] for s in g["stats"])),
                          "digest": digest(g, output)}
            all_ok = all_ok and not bad
        nat._dll.cudapoa_b200_debug_set_schedule(0)
        report["cases"].append(row)
        print(json.dumps(row), flush=True)
    report["all_equal"] = all_ok
    with open(out_path, "w") as f:
        json.dump(report, f, indent=1)
    print("all equal:", all_ok)
    return 0 if all_ok else 1


if __name__ == "__main__":
    sys.exit(main())
