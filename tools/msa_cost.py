"""Development helper (GPU box): device time of a cfg2-shaped batch with consensus and with MSA output."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch  # noqa: E402

lib = _cabi.native()
dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
dll.cudapoa_b200_bench_relaunch.restype = ctypes.c_double
dll.cudapoa_b200_bench_relaunch.argtypes = [ctypes.c_void_p, ctypes.c_int32]
dll.cudapoa_b200_bench_relaunch_phases.restype = ctypes.c_int32
dll.cudapoa_b200_bench_relaunch_phases.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 9472
ws, _ = bench.make_windows(n, 1024, 0)
for out in ("consensus", "msa"):
    b = CudaPoaBatch(30, 640, 100 << 30, output_type=out, cuda_banded_alignment=True, alignment_band_width=256, lib=lib)
    b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
    b.generate_poa()
    b.fetch(out)
    dll.cudapoa_b200_bench_relaunch(b._handle, 1)
    ms = dll.cudapoa_b200_bench_relaunch(b._handle, 3) / 3
    f, g = ctypes.c_double(0), ctypes.c_double(0)
    dll.cudapoa_b200_bench_relaunch_phases(b._handle, ctypes.byref(f), ctypes.byref(g))
    print("%-9s %d windows: %.1f ms per batch (fill %.1f, graph %.1f) = %.0f windows/s" % (out, n, ms, f.value, g.value, n / ms * 1e3), flush=True)
    b.close()
