#!/usr/bin/env python
"""Development helper (no GPU): a wider sweep of windows through the SIMT emulator (tests/emu) against the CPU oracle than
the test suite runs -- full cfg 2 windows, noisy windows with many reads (rows with 2..6 predecessors, far predecessors),
long rebased windows, full alignment.  `python tools/emu_check.py [quick]`"""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "emu")]
import oraclelib  # noqa: E402
import parity  # noqa: E402
import build as emu_build  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.windows import generate_windows  # noqa: E402

os.environ["POA_EMU_RCP_TABLE"] = oraclelib.RCP_TABLE
lib = _cabi.CabiLibrary(emu_build.build(), "cudapoa_b200_")
set_schedule = lib._dll.cudapoa_b200_debug_set_schedule
set_schedule.argtypes = [ctypes.c_int32]
quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
scale = int(os.environ.get("EMU_SCALE", "1"))

CASES = [
    ("cfg2 x3", dict(n=2 if quick else 3, start_index=100, template_length=500, n_reads=30),
     dict(max_seq=640, max_reads=30, band=256, banded=True, output="consensus"), 2),
    ("noisy 300x20", dict(n=2 if quick else 4, start_index=4100, template_length=300, n_reads=20, read_error=0.25, backbone_error=0.15),
     dict(max_seq=448, max_reads=20, band=256, banded=True, output="consensus"), 2),
    ("noisy 200x12 single launch", dict(n=3, start_index=4200, template_length=200, n_reads=12, read_error=0.3, backbone_error=0.2),
     dict(max_seq=320, max_reads=12, band=256, banded=True, output="consensus"), 1),
    ("long 1800x6 rebased", dict(n=1, start_index=7100, template_length=1800, n_reads=6, read_error=0.15),
     dict(max_seq=2400, max_reads=6, band=256, banded=True, output="consensus"), 2),
    ("full 400x8", dict(n=2, start_index=8100, template_length=400, n_reads=8, read_error=0.15),
     dict(max_seq=512, max_reads=8, band=256, banded=False, output="consensus"), 2),
    ("short 90x6", dict(n=3, start_index=9500, template_length=90, n_reads=6, read_error=0.15),
     dict(max_seq=256, max_reads=6, band=256, banded=True, output="consensus"), 2),
    ("full 700x5 (3 chunks)", dict(n=2, start_index=9600, template_length=700, n_reads=5, read_error=0.12),
     dict(max_seq=768, max_reads=5, band=256, banded=False, output="consensus"), 2),
    ("mixed 60..620 x 8", dict(n=6, start_index=9700, template_length=(60, 620), n_reads=8, read_error=0.15),
     dict(max_seq=704, max_reads=8, band=256, banded=True, output="consensus"), 2),
    ("msa noisy 250x10", dict(n=2, start_index=9100, template_length=250, n_reads=10, read_error=0.2),
     dict(max_seq=384, max_reads=10, band=256, banded=True, output="msa"), 2),
]
bad_total = 0
for name, gen, eng, schedule in CASES:
    gen = dict(gen)
    n = gen.pop("n") * scale
    ws = generate_windows(n, **gen)
    set_schedule(schedule)
    t = time.time()
    out = parity.run_abi(lib, ws, want_graphs=False, max_mem=1 << 30, **eng)
    o = parity.run_oracle(ws, want_graphs=False, **eng)
    bad = parity.compare(out, o, eng["output"])
    exact = sum(s["exact_path"] for s in out["stats"])
    print("%-28s windows %d  mismatching %s  exact-path alignments %d  status %s  %.1fs" % (name, n, bad, exact, sorted(set(out["status"])), time.time() - t), flush=True)
    bad_total += len(bad)
set_schedule(0)
sys.exit(1 if bad_total else 0)
