"""Development helper (GPU box): device time of one lone batch under the single-launch and the phase-split schedule
for several batch sizes (cfg2 windows)."""
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
sizes = [int(x) for x in sys.argv[1:]] or [148, 592, 1184, 2368, 4736, 9472]
ws_all, _ = bench.make_windows("cfg2", max(sizes), 4096, 0)
batch = CudaPoaBatch(30, 640, 40 << 30, output_type="consensus", cuda_banded_alignment=True, alignment_band_width=256,
                     lib=lib)
for n in sizes:
    ws = ws_all.slice(0, n)
    row = []
    for mode in (1, 2):
        lib._dll.cudapoa_b200_debug_set_schedule(mode)
        batch.reset()
        batch.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
        batch.generate_poa()
        batch.fetch("consensus")
        dll.cudapoa_b200_bench_relaunch(batch._handle, 2)
        ms = dll.cudapoa_b200_bench_relaunch(batch._handle, 5) / 5
        row.append(ms)
        import time
        t0 = time.perf_counter()
        for _ in range(3):
            batch.reset()
            batch.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
            t1 = time.perf_counter()
            batch.generate_poa()
            t2 = time.perf_counter()
            batch.fetch("consensus")
        t3 = time.perf_counter()
        row.append((t3 - t0) / 3 * 1e3)
        row.append((t2 - t1) * 1e3)
    print("windows %6d: single launch %8.2f ms (wall %6.1f, enqueue %5.2f), phase-split %8.2f ms (wall %6.1f, enqueue %5.2f)"
          % (n, row[0], row[1], row[2], row[3], row[4], row[5]), flush=True)
