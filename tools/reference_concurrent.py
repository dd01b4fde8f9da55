"""Run ON A B200.  The unmodified reference cudapoa (oracle/_ref/libcudapoa_ref.so, sm_100a build with its own flags)
timed the way its own benchmark drives it -- K concurrent batches on K streams from K host threads
(cudapoa/benchmarks/multi_batch.hpp:43-56,157-170) -- for K in {1, 2, 4, 8, 16}, memory budget = a share of free memory
per batch, on BASELINE.json's cfg 2 windows (banded 256, consensus).  Prints one JSON line per K and a summary; the
best K is "the kernel to beat" quoted in BASELINE.md.  With --engine native the same driver runs this repository's
library for an equal-footing comparison.

Usage: python tools/reference_concurrent.py [--windows N] [--ks 1,2,4,8,16] [--engine ref|native] [--config cfg2]
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402  (streams and memory info only)

from claragenomicsanalysis_b200 import _cabi  # noqa: E402
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch  # noqa: E402
from claragenomicsanalysis_b200.windows import config_windows  # noqa: E402

SHAPES = {"cfg2": dict(max_seq=640, max_reads=30, output="consensus"),
          "cfg3": dict(max_seq=640, max_reads=30, output="msa"),
          "cfg4": dict(max_seq=6400, max_reads=100, output="consensus"),
          "cfg5b": dict(max_seq=2560, max_reads=60, output="consensus")}


def run_k(lib, ws, k, shape, mem_per_batch, banded=True):
    """All windows of `ws` through k batches (k threads, k streams); each thread takes chunks of its batch's capacity
    from a shared cursor.  Returns (seconds, windows done, rounds, capacity)."""
    streams = [torch.cuda.Stream() for _ in range(k)]
    batches = [CudaPoaBatch(shape["max_reads"], shape["max_seq"], mem_per_batch, output_type=shape["output"],
                            stream=streams[i], cuda_banded_alignment=banded, alignment_band_width=256, lib=lib)
               for i in range(k)]
    lock = threading.Lock()
    state = dict(next=0, done=0, rounds=0, cap=0)
    n = len(ws)
    chunk_hint = [n]  # windows offered to a batch per round; shrinks to the capacity once a batch reports full

    def worker(b):
        # as multi_batch.hpp:157-170: the shared window list is consumed under a mutex (staging is serialised, the
        # device work and the result fetch of the K batches overlap)
        while True:
            with lock:
                start = state["next"]
                if start >= n:
                    return
                part = ws.slice(start, min(n, start + chunk_hint[0]))
                b.reset()
                rc, added = b.add_poa_groups_flat(part.bases, part.read_lengths, part.window_n_reads)
                if added == 0:
                    raise RuntimeError("batch accepted nothing (status %d)" % rc)
                if rc == 1:  # exceeded_maximum_poas: the batch is full, this is its capacity
                    chunk_hint[0] = added
                state["next"] = start + added
                state["cap"] = max(state["cap"], added)
            b.generate_poa()
            b.fetch(shape["output"])
            with lock:
                state["done"] += added
                state["rounds"] += 1

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    threads = [threading.Thread(target=worker, args=(b,)) for b in batches]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    for b in batches:
        b.close()
    return dt, state["done"], state["rounds"], state["cap"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--windows", type=int, default=16384)
    ap.add_argument("--distinct", type=int, default=2048)
    ap.add_argument("--ks", default="1,2,4,8,16")
    ap.add_argument("--engine", default="ref", choices=["ref", "native"])
    ap.add_argument("--config", default="cfg2", choices=sorted(SHAPES))
    ap.add_argument("--mem-share", type=float, default=0.8, help="share of free memory split among the K batches")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--full", action="store_true", help="full alignment instead of the static band")
    args = ap.parse_args()
    if args.engine == "ref":
        import parity

        lib = parity.ref_lib()
    else:
        lib = _cabi.native()
    shape = SHAPES[args.config]
    base = config_windows(args.config, min(args.distinct, args.windows))
    ws = base.tile(args.windows)
    best = None
    for k in [int(x) for x in args.ks.split(",")]:
        free_b, _ = torch.cuda.mem_get_info(0)
        mem = int(free_b * args.mem_share / k)
        times = []
        for _ in range(args.reps):
            dt, done, rounds, cap = run_k(lib, ws, k, shape, mem, banded=not args.full)
            times.append(dt)
        dt = min(times)
        line = dict(engine=args.engine, config=args.config, full_alignment=args.full, k=k, windows=done, seconds=round(dt, 4),
                    windows_per_s=round(done / dt, 1), rounds=rounds, windows_per_batch=cap,
                    one_warp_blocks_per_wave=148 * 32,
                    waves_per_batch=round(cap / (148 * 32.0), 2), mem_per_batch_gb=round(mem / 2**30, 2))
        print(json.dumps(line), flush=True)
        if best is None or line["windows_per_s"] > best["windows_per_s"]:
            best = line
    print(json.dumps(dict(best=best)), flush=True)


if __name__ == "__main__":
    main()
