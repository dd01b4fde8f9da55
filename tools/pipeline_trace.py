"""Development helper (GPU box): one traced run of the window pipeline (CUDAPOA_B200_PIPELINE_TRACE=1): where the wall
time of an end-to-end step goes -- per chunk, when its staging, launch and collection started and ended."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from claragenomicsanalysis_b200.pipeline import WindowPipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 18944
ws, base = bench.make_windows("cfg2", n, 4096, 0)
for spec in sys.argv[2:] or ["8,1184,6"]:
    workers, chunk, mem = (int(x) for x in spec.split(","))
    pipe = WindowPipeline(30, 640, mem << 30, devices=(0,), workers_per_device=workers, output_type="consensus",
                          cuda_banded_alignment=True, alignment_band_width=256)
    pipe.run(ws, chunk)
    pipe.run(ws, chunk)
    os.environ["CUDAPOA_B200_PIPELINE_TRACE"] = "1"
    t0 = time.perf_counter()
    pipe.run(ws, chunk)
    dt = time.perf_counter() - t0
    os.environ["CUDAPOA_B200_PIPELINE_TRACE"] = "0"
    print("workers %d chunk %d: %.1f ms wall, %.0f windows/s" % (workers, chunk, dt * 1e3, n / dt), flush=True)
    pipe.close()
