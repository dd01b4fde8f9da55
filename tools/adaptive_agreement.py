"""GPU box: agreement numbers of adaptive banding (an extension with no reference counterpart, hence
"unpinned").  For structural-variant windows (reads that leave the diagonal by more than a static band follows) and for
ordinary cfg 2 windows: consensus of the static band, of adaptive banding and of the full alignment through the engine,
and SPOA (3rdparty/spoa, full alignment) on the host cores; reported as fractions of windows whose consensus is
identical.  One JSON object on stdout.

    python tools/adaptive_agreement.py [n_structural] [n_cfg2]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
import oraclelib  # noqa: E402
from claragenomicsanalysis_b200.windows import config_windows_parallel  # noqa: E402


def engine_consensus(lib, ws, max_seq, max_reads, band, banded, mem=40 << 30):
    from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

    out = []
    stats = {"widened": 0, "exact_path": 0, "reads_aligned": 0}
    t0 = time.perf_counter()
    batch = CudaPoaBatch(max_reads, max_seq, mem, output_type="consensus", device_id=0, cuda_banded_alignment=banded,
                         alignment_band_width=band, lib=lib)
    try:
        at = 0
        while at < len(ws):
            part = ws.slice(at, min(len(ws), at + batch.max_poas))
            batch.reset()
            rc, added = batch.add_poa_groups_flat(part.bases, part.read_lengths, part.window_n_reads)
            if added <= 0:
                raise RuntimeError("window %d does not fit a batch (status %d)" % (at, rc))
            batch.generate_poa()
            cons, _, status = batch.get_consensus()
            out.extend(c if s == 0 else None for c, s in zip(cons[:added], status[:added]))
            st = batch.stats()
            for k in stats:
                stats[k] += st[k]
            at += added
    finally:
        batch.close()
    return out, stats, time.perf_counter() - t0


def agreement(a, b):
    return sum(1 for x, y in zip(a, b) if x is not None and x == y) / max(len(a), 1)


def study(lib, name, ws, max_seq, max_reads, band):
    static, st_s, t_s = engine_consensus(lib, ws, max_seq, max_reads, band, True)
    adaptive, st_a, t_a = engine_consensus(lib, ws, max_seq, max_reads, band, "adaptive")
    full, st_f, t_f = engine_consensus(lib, ws, max_seq, max_reads, band, False)
    line = {"windows": len(ws), "band": band, "batch_size": "BatchSize(%d, %d, %d)" % (max_seq, max_reads, band),
            "static_vs_full": agreement(static, full), "adaptive_vs_full": agreement(adaptive, full),
            "adaptive_vs_static": agreement(adaptive, static),
            "widened_alignments": st_a["widened"], "alignments": st_a["reads_aligned"],
            "exact_path_alignments_adaptive": st_a["exact_path"],
            "seconds": {"static": round(t_s, 2), "adaptive": round(t_a, 2), "full": round(t_f, 2)}}
    if oraclelib.have_ref_spoa():
        r = oraclelib.SpoaRun(ws, match=8, mismatch=-6, gap=-8, msa=False, threads=os.cpu_count() or 1)
        spoa = [r.consensus(w) for w in range(len(ws))]
        line.update({"static_vs_spoa": agreement(static, spoa), "adaptive_vs_spoa": agreement(adaptive, spoa),
                     "full_vs_spoa": agreement(full, spoa), "seconds_spoa": round(r.seconds, 2)})
    print(name, json.dumps(line), file=sys.stderr, flush=True)
    return line


def main():
    n_struct = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    n_cfg2 = int(sys.argv[2]) if len(sys.argv) > 2 else 20480
    cfg2 = config_windows_parallel("cfg2", n_cfg2, start_index=500000)  # forks: before CUDA is initialised
    from claragenomicsanalysis_b200 import _cabi

    lib = _cabi.native()
    structural = cases.structural_windows(n_struct, start=300000)
    out = {"structural_band128": study(lib, "structural_band128", structural, 1024, 12, 128),
           "structural_band256": study(lib, "structural_band256", structural, 1024, 12, 256),
           "cfg2_band256": study(lib, "cfg2_band256", cfg2, 640, 30, 256),
           "note": "fractions of windows with identical consensus; full = full (unbanded) alignment through the engine, "
                   "the reference's own default mode; adaptive banding has no reference implementation (parity unpinned)"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
