"""Development helper (GPU box): the window pipeline run repeatedly, every run compared window by window with the first
(results must not depend on how batches overlap on the device)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from claragenomicsanalysis_b200.pipeline import WindowPipeline  # noqa: E402

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n = int(sys.argv[2]) if len(sys.argv) > 2 else 9472
workers = int(sys.argv[3]) if len(sys.argv) > 3 else 3
chunk = int(sys.argv[4]) if len(sys.argv) > 4 else 2368
ws, base = bench.make_windows(n, 1024, 0)
pipe = WindowPipeline(30, 640, 8 << 30, devices=(0,), workers_per_device=workers, output_type="consensus",
                      cuda_banded_alignment=True, alignment_band_width=256)
pipe.run(ws, chunk)
ref = [pipe.consensus(i) for i in range(n)]
ref_st = pipe.statuses()
print("first run: statuses ok", sum(1 for s in ref_st if s == 0), "of", n)
# windows repeat with period 1024: equal inputs must give equal outputs within a run as well
bad0 = [i for i in range(n) if ref[i] != ref[i % 1024] or ref_st[i] != ref_st[i % 1024]]
print("first run: windows differing from their first copy:", bad0[:20], len(bad0))
for r in range(runs):
    pipe.run(ws, chunk)
    st = pipe.statuses()
    bad = [i for i in range(n) if st[i] != ref_st[i % 1024] or pipe.consensus(i) != ref[i % 1024]]
    print("run", r, "differing windows:", len(bad), bad[:12], [st[i] for i in bad[:12]])
    for i in bad[:3]:
        a, b = pipe.consensus(i), ref[i % 1024]
        print("   window", i, "len", len(a[0]), len(b[0]), "first diff",
              next((k for k in range(min(len(a[0]), len(b[0]))) if a[0][k] != b[0][k]), None),
              "coverage diffs (pos, got, want):", [(k, x, y) for k, (x, y) in enumerate(zip(a[1], b[1])) if x != y][:10],
              "n", sum(1 for x, y in zip(a[1], b[1]) if x != y))
pipe.close()

# the same windows through ONE large batch (the bench's resident leg), several times
if len(sys.argv) > 5:
    from claragenomicsanalysis_b200 import _cabi
    from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch
    big = int(sys.argv[5])
    ws2, _ = bench.make_windows(big, 1024, 0)
    batch = CudaPoaBatch(30, 640, 70 << 30, output_type="consensus", device_id=0, cuda_banded_alignment=True,
                         alignment_band_width=256, lib=_cabi.native())
    for r in range(int(sys.argv[6]) if len(sys.argv) > 6 else 4):
        batch.reset()
        batch.add_poa_groups_flat(ws2.bases, ws2.read_lengths, ws2.window_n_reads)
        batch.generate_poa()
        cons, cov, st = batch.get_consensus()
        bad = [i for i in range(big) if st[i] != ref_st[i % 1024] or (cons[i], list(cov[i])) != ref[i % 1024]]
        print("big batch run", r, "differing windows:", len(bad), bad[:12], [st[i] for i in bad[:12]])
        for i in bad[:3]:
            a, b = cons[i], ref[i % 1024][0]
            print("   window", i, "len", len(a), len(b), "first diff",
                  next((k for k in range(min(len(a), len(b))) if a[k] != b[k]), None),
                  "cov equal", list(cov[i]) == ref[i % 1024][1])
