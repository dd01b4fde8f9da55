"""Development helper: run the first `n_reads` reads of window `w` of a parity case through the phase-split schedule and
dump the window's arena and score area (works on the GPU library and on the emulator build: POA_DUMP_EMU=1)."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch  # noqa: E402
from claragenomicsanalysis_b200.windows import WindowSet  # noqa: E402

name, w, n_reads, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
if os.environ.get("POA_DUMP_EMU"):
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    import build as emu_build
    import oraclelib

    os.environ["POA_EMU_RCP_TABLE"] = oraclelib.RCP_TABLE
    lib = _cabi.CabiLibrary(emu_build.build(), "cudapoa_b200_")
else:
    lib = _cabi.native()
lib._dll.cudapoa_b200_debug_set_schedule(2)
c = cases.CASES[name]
reads = cases.case_windows(name).window(w)[:n_reads]
ws = WindowSet.from_lists([reads])
eng = c["engine"]
b = CudaPoaBatch(eng["max_reads"], eng["max_seq"], 1 << 30, output_type=eng["output"], cuda_banded_alignment=eng["banded"],
                 alignment_band_width=eng["band"], lib=lib)
b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
b.generate_poa()
cons, cov, st = b.get_consensus()
f = lib._dll.cudapoa_b200_debug_read_window
f.restype = ctypes.c_int64
f.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int64]
res = {"consensus": np.frombuffer(cons[0].encode(), dtype=np.uint8)}
for region, key in ((0, "arena"), (1, "scores")):
    size = f(b._handle, 0, region, None, 0)
    buf = np.zeros(size, dtype=np.uint8)
    f(b._handle, 0, region, buf.ctypes.data, size)
    res[key] = buf
np.savez_compressed(out, **res)
print("dumped", name, w, n_reads, {k: v.shape for k, v in res.items()}, "status", st, "len", len(cons[0]))
