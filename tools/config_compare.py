"""Development helper (GPU box): every BASELINE.json configuration through the native engine and through the
reference cudapoa build (oracle/_ref) behind the same C ABI, on the same device -- throughput and output equality.
Reads nothing from /root/reference.  `python tools/config_compare.py [cfg ...]`"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity  # noqa: E402
from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch  # noqa: E402
from claragenomicsanalysis_b200.windows import config_windows  # noqa: E402

CASES = {
    # name: (generator config, windows, max_seq, max_reads, output, banded, memory, windows per round)
    "cfg2": ("cfg2", 2368, 640, 30, "consensus", True, 16 << 30, 1184),
    "cfg3": ("cfg3", 2368, 640, 30, "msa", True, 16 << 30, 1184),
    "cfg2_full": ("cfg2", 592, 640, 30, "consensus", False, 16 << 30, 296),
    "cfg4": ("cfg4", 64, 6400, 100, "consensus", True, 32 << 30, 16),
    "cfg4_msa": ("cfg4", 32, 6400, 100, "msa", True, 32 << 30, 8),
    "cfg5b": ("cfg5b", 1024, 2560, 60, "consensus", True, 32 << 30, 256),
    "cfg5b_big": ("cfg5b", 4096, 2560, 60, "consensus", True, 100 << 30, 2048),
    "cfg5b_huge": ("cfg5b", 8192, 2560, 60, "consensus", True, 130 << 30, 8192),
}


def run(lib, nat_lib, ws, max_seq, max_reads, output, banded, mem, chunk):
    b = CudaPoaBatch(max_reads, max_seq, mem, output_type=output, cuda_banded_alignment=banded,
                     alignment_band_width=256, lib=lib)
    digests = []
    t_total = 0.0
    start = 0
    n = len(ws)
    rounds = 0
    while start < n:
        part = ws.slice(start, min(n, start + chunk))
        t0 = time.perf_counter()
        rc, added = b.add_poa_groups_flat(part.bases, part.read_lengths, part.window_n_reads)
        if added != len(part):
            raise RuntimeError("round of %d windows: %d accepted, rc=%d" % (len(part), added, rc))
        b.generate_poa()
        b.fetch(output)
        t_total += time.perf_counter() - t0
        digests.append(b.result_digest())
        st = b.stats() if lib is nat_lib else None
        b.reset()
        start += added
        rounds += 1
    b.close()
    return t_total, digests, rounds, st


def main():
    names = sys.argv[1:] or list(CASES)
    ref = parity.ref_lib()
    nat = _cabi.native()
    for name in names:
        cfg, n, max_seq, max_reads, output, banded, mem, chunk = CASES[name]
        ws = config_windows(cfg, n)
        out = {"case": name, "windows": n}
        for label, lib in (("native", nat), ("reference", ref)):
            if label == "reference" and os.environ.get("COMPARE_NATIVE_ONLY"):
                continue
            try:
                run(lib, nat, ws.slice(0, min(n, 16)), max_seq, max_reads, output, banded, mem, chunk)  # warm-up
                t, dg, rounds, st = run(lib, nat, ws, max_seq, max_reads, output, banded, mem, chunk)
                out[label] = {"windows_per_s": round(n / t, 1), "seconds": round(t, 3), "rounds": rounds}
                out[label + "_digest"] = dg
                if st is not None:
                    out["native_exact_path"] = st.get("exact_path")
            except Exception as e:
                out[label] = {"error": str(e)[:200]}
        if "native_digest" in out and "reference_digest" in out:
            out["equal"] = out["native_digest"] == out["reference_digest"]
        out.pop("native_digest", None)
        out.pop("reference_digest", None)
        if "windows_per_s" in out.get("native", {}) and "windows_per_s" in out.get("reference", {}):
            out["speedup"] = round(out["native"]["windows_per_s"] / out["reference"]["windows_per_s"], 2)
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
