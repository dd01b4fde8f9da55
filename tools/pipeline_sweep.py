"""Development helper (GPU box): end-to-end windows/s of the window pipeline for several (workers, chunk) settings."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from claragenomicsanalysis_b200.pipeline import WindowPipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 37888
ws, base = bench.make_windows("cfg2", n, 4096, 0)
for spec in sys.argv[2:]:
    workers, chunk, mem = (int(x) for x in spec.split(","))
    pipe = WindowPipeline(30, 640, mem << 30, devices=(0,), workers_per_device=workers, output_type="consensus",
                          cuda_banded_alignment=True, alignment_band_width=256)
    pipe.run(ws, chunk)
    t0 = time.perf_counter()
    for _ in range(3):
        pipe.run(ws, chunk)
    dt = (time.perf_counter() - t0) / 3
    print("workers %d chunk %d mem %d GB: %.0f windows/s" % (workers, chunk, mem, n / dt), flush=True)
    pipe.close()
