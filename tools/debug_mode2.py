"""Development helper (GPU box): run parity cases through the phase-split schedule (mode 2) and the single launch
(mode 1) and report, per case, which windows differ from the CPU oracle."""
import sys
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity  # noqa: E402
import cases  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402

lib = _cabi.native()
names = sys.argv[2:] or ["banded16_short_reads", "banded16_weighted", "cfg2_banded16_consensus", "banded16_noisy"]
nmax = int(sys.argv[1]) if len(sys.argv) > 1 else 4
for name in names:
    c = cases.CASES[name]
    ws = cases.case_windows(name).slice(0, nmax)
    eng = c["engine"]
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    for mode in (1, 2):
        lib._dll.cudapoa_b200_debug_set_schedule(mode)
        r = parity.run_abi(lib, ws, want_graphs=True, max_mem=2 << 30, **eng)
        bad = parity.compare(r, o, eng["output"])
        gbad = [w for w in range(len(ws)) if r["status"][w] == 0 and r["graphs"][w] != o["graphs"][w]]
        st = r["stats"][0]
        print(name, "mode", mode, "status", sorted(set(r["status"])), "bad", bad, "graphbad", gbad, "exact", st["exact_path"],
              "launches", st["kernel_launches"], flush=True)
        if bad and eng["output"] == "consensus":
            w = bad[0]
            a, b = r["consensus"][w], o["consensus"][w]
            k = next((i for i in range(min(len(a), len(b))) if a[i] != b[i]), min(len(a), len(b)))
            print("   window", w, "len", len(a), len(b), "first diff at", k, "nodes", len(r["graphs"][w][0]), len(o["graphs"][w][0]))
lib._dll.cudapoa_b200_debug_set_schedule(0)
