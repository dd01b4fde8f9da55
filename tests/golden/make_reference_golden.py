"""Run ON A B200 (needs oracle/_ref/libcudapoa_ref.so, built from /root/reference by oracle/Makefile).

For every case of tests/cases.py: run the UNMODIFIED reference cudapoa (sm_100a build, its own flags), store its
per-window outputs as tests/golden/reference_<case>.json, and report how the CPU oracle and the native engine
compare.  The JSON files are the "outputs of the reference itself" that pin the oracle.
Usage: python tests/golden/make_reference_golden.py [case ...]
"""
import json
import os
import shutil
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity  # noqa: E402
from cases import CASES, case_windows  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402


def main():
    names = sys.argv[1:] or list(CASES)
    out_dir = os.path.join(ROOT, "tests", "golden")
    scratch = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(scratch, exist_ok=True)
    ref = parity.ref_lib()
    try:
        native = _cabi.native()
    except Exception as e:  # the native engine is optional for golden generation
        print("native library unavailable:", e)
        native = None
    summary = {}
    for name in names:
        case = CASES[name]
        ws = case_windows(name)
        eng = case["engine"]
        output = eng["output"]
        t0 = time.time()
        r = parity.run_abi(ref, ws, want_graphs=True, **eng)
        t_ref = time.time() - t0
        doc = dict(case=name, n_windows=len(ws), engine=eng, gen=case["gen"], start=case.get("start", 0),
                   implementation="genomeworks-reference sm_100a -use_fast_math",
                   status=[int(s) for s in r["status"]], seq_status=r["seq_status"],
                   digest=parity.digests(r, output))
        if output == "consensus":
            doc["consensus"] = r["consensus"]
            doc["coverage"] = [list(map(int, c)) for c in r["coverage"]]
        else:
            doc["msa"] = r["msa"]
        doc["graph_nodes"] = [g[0].decode("latin1") for g in r["graphs"][:4]]
        doc["graph_node_count"] = [len(g[0]) for g in r["graphs"]]
        doc["graph_edges"] = [[list(e) for e in g[1]] for g in r["graphs"][:4]]
        doc["graph_edge_count"] = [len(g[1]) for g in r["graphs"]]
        path = os.path.join(out_dir, "reference_%s.json" % name)
        with open(path, "w") as f:
            json.dump(doc, f, separators=(",", ":"))
        shutil.copy(path, scratch)
        line = dict(ref_seconds=round(t_ref, 2), statuses=sorted(set(doc["status"])))
        # oracle vs reference
        t0 = time.time()
        o = parity.run_oracle(ws, want_graphs=True, **eng)
        line["oracle_seconds"] = round(time.time() - t0, 2)
        line["oracle_mismatch"] = parity.compare(r, o, output)
        line["oracle_graph_mismatch"] = [w for w in range(len(ws))
                                         if r["status"][w] == 0 and (r["graphs"][w][0] != o["graphs"][w][0] or
                                                                     r["graphs"][w][1] != o["graphs"][w][1])]
        line["oracle_rcp_misses"] = o["rcp_misses"]
        if native is not None:
            t0 = time.time()
            n = parity.run_abi(native, ws, want_graphs=True, **eng)
            line["native_seconds"] = round(time.time() - t0, 2)
            line["native_mismatch"] = parity.compare(r, n, output)
            line["native_graph_mismatch"] = [w for w in range(len(ws))
                                             if r["status"][w] == 0 and (r["graphs"][w][0] != n["graphs"][w][0] or
                                                                         r["graphs"][w][1] != n["graphs"][w][1])]
            line["native_seq_status_equal"] = (n["seq_status"] == r["seq_status"])
            line["native_stats"] = n.get("stats", [None])[0]
        summary[name] = line
        print(name, json.dumps(line), flush=True)
    with open(os.path.join(scratch, "summary.json"), "w") as f:
        json.dump(summary, f, indent=1)


if __name__ == "__main__":
    main()
