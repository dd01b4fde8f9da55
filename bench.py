#!/usr/bin/env python
"""Benchmark of the B200 cudapoa engine on BASELINE.json's metric: POA windows/s (+ GCUPS).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

Headline workload at every N: BASELINE.json configs[1] -- Racon-style polishing windows, 500 bp template x 30 ONT-like
reads, static band 256, int16 scores, consensus output (SURVEY.md 8(d) "cfg 2").  One "step" = one batch of `--windows`
such windows per GPU through the hot path.  Windows are independent: each rank owns its own windows (weak scaling), no
collective on the data path; torch.distributed is used for the barrier and the max-over-ranks only.

Native arm, per run:
  value      windows/s with the batch's inputs already resident in HBM: K passes of the batch's kernels timed with CUDA
             events on the batch's stream (cudapoa_b200_bench_relaunch_timed), max over ranks.
  e2e        windows/s through the C ABI with HOST buffers every step: the window pipeline (cudapoa_b200_pipeline_run)
             drives `--workers` Batch objects, each reset -> add_poa_group (host staging) -> generate_poa (H2D + kernels)
             -> get_consensus (D2H + decode) per chunk of windows; wall clock between device syncs.
  roofline   the binding view: GCUPS of the dominant kernel against the s16x2 DPX issue rate measured
             in this run; `roofline_hbm` is the secondary HBM view on the algorithmic bytes.
  configs    the other named shapes of BASELINE.json (cfg 3 MSA, cfg 4 long windows, cfg 5B mixed sizes), a few steps
             each: resident windows/s, end-to-end windows/s through one Batch, GCUPS.
  cpu_baseline  the reference repository's CPU implementation of the same job (SPOA, 3rdparty/spoa, built unmodified
             into oracle/_ref/libspoa_ref.so) on all host cores over a bounded sample, plus consensus agreement rate.

Reference arm (--impl reference): SPOA on the host cores, same metric/config/unit, bounded sample per step.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "poa_windows_per_second"
UNIT = "windows/s"
WORKLOAD = "cfg2: 500 bp x 30 ONT-like reads, banded(256) int16 consensus, scores 8/-6/-8"
ENGINE = dict(max_seq=640, max_reads=30, band=256)
# the other named shapes: generator name, BatchSize, output, windows per GPU and step, steps
OTHER_CONFIGS = {
    "cfg3": dict(gen="cfg3", max_seq=640, max_reads=30, output="msa", windows=4736, steps=2,
                 workload="cfg3: cfg2 windows, MSA output, static band 256"),
    "cfg4": dict(gen="cfg4", max_seq=6400, max_reads=100, output="consensus", windows=592, steps=1,
                 workload="cfg4: 2-5 kb x 100 reads, band 256, int32 scores in the reference"),
    "cfg5b": dict(gen="cfg5b", max_seq=2560, max_reads=60, output="consensus", windows=2368, steps=2,
                  workload="cfg5b: templates 250/500/1000/2000 bp (p .2/.5/.2/.1) x 10..60 reads in one batch"),
    "cfg2_full": dict(gen="cfg2", max_seq=640, max_reads=30, output="consensus", windows=9472, steps=2, banded=False,
                      workload="cfg2 windows, full alignment (the reference's default mode; cudapoa/benchmarks/"
                               "single_batch.hpp:48-50): packed fill in column chunks of 256 cells"),
}


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.rows = []
        self.stamps = []
        self.proc = None
        self.thread = None
        self.t_begin = 0.0

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 8:
                self.rows.append(parts)
                self.stamps.append(time.time())

    def mark(self):
        """The timed region starts now (the sampler itself is started earlier: nvidia-smi takes a second or more to
        produce its first line on a box with many GPUs)."""
        self.t_begin = time.time()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t_end = time.time()
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = [r for r, t in zip(self.rows, self.stamps) if self.t_begin <= t <= t_end + 0.05]
        window = "timed region"
        if not rows:  # region shorter than the sampling period: the warm-up steps just before ran the same load
            rows = [r for r, t in zip(self.rows, self.stamps) if t >= self.t_begin - 2.0]
            window = "last 2 s of warm-up + timed region (same load; no sample fell inside the timed region)"
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in rows if r[3].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "window": window}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_windows(name, n_windows, distinct, start_index):
    """`n_windows` windows of a named shape: min(distinct, n_windows) generated (process pool), tiled to the count the
    way the reference tiles its input file.  Call before CUDA is initialised (the pool forks)."""
    from claragenomicsanalysis_b200.windows import config_windows_parallel

    base = config_windows_parallel(name, min(distinct, n_windows), start_index=start_index)
    return base.tile(n_windows), base


def algorithmic_bytes(ws, total_consensus_chars):
    """bases + 32 B descriptor per window + 2 B per read length + consensus (1 B) and coverage (2 B)
    per consensus base.  Weights are not transferred when every weight is 1."""
    return int(ws.n_bases + 32 * len(ws) + 2 * ws.n_reads + 3 * total_consensus_chars)


def spoa_run(ws, threads=0):
    import oraclelib

    return oraclelib.SpoaRun(ws, match=8, mismatch=-6, gap=-8, msa=False, threads=threads)


def ncu_traffic():
    """DRAM bytes per window and step of each kernel, from the committed ncu launch list (profiles/ncu_traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum summed over the launches of one step, divided by its windows)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


class _StdoutToStderr:
    """fd-level redirect: native code that logs on stdout must not break the one-JSON-line contract."""

    def __enter__(self):
        sys.stdout.flush()
        self._saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self._saved, 1)
        os.close(self._saved)
        return False


def guarded(fn, *a, **kw):
    """Reporting legs must never cost the run its JSON line: their failure is recorded, not raised."""
    try:
        return fn(*a, **kw)
    except Exception as e:  # noqa: BLE001
        return {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}


def run_reference(args, rank, world):
    """SPOA (the reference repository's CPU POA) on all host cores; rank 0 only."""
    if rank != 0:
        return
    import oraclelib

    if not oraclelib.have_ref_spoa():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libspoa_ref.so was not built"}))
        return
    cores = host_cores()
    probe, _ = make_windows("cfg2", max(2 * cores, 8), 4096, 0)
    r = spoa_run(probe, cores)
    rate = len(probe) / max(r.seconds, 1e-6)
    budget = min(10.0, 150.0 / max(args.steps + args.warmup, 1))
    n = max(cores, int(rate * budget))
    ws, _ = make_windows("cfg2", n, 65536, 0)
    for _ in range(args.warmup):
        spoa_run(ws, cores)
    t = 0.0
    cells = 0
    for _ in range(args.steps):
        r = spoa_run(ws, cores)
        t += r.seconds
        cells += r.cells
    value = n * args.steps / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "windows_per_step": n, "engine": "SPOA v3.0.0 kNW linear gaps, full (unbanded)",
                   "same_config_as_native_arm": False,
                   "note": "the reference repository's CPU implementation of the job (full alignment, its only mode) on "
                           "a bounded sample per step; the GPU arm runs the banded(256) configuration on 18 944 windows"},
        "gcups": cells / t / 1e9,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "%d cfg2 windows per step x %d steps, SPOA on %d threads" % (n, args.steps, cores)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def cpu_baseline(args, base, batch):
    """SPOA on all host cores over a bounded sample of the same windows + consensus agreement with the engine."""
    import oraclelib

    if not oraclelib.have_ref_spoa():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "libspoa_ref.so not built"}
    cores = host_cores()
    probe = base.slice(0, max(2 * cores, 8))
    r = spoa_run(probe, cores)
    rate = len(probe) / max(r.seconds, 1e-6)
    # bounded by time AND by what one batch holds (the agreement check runs the same windows through the engine)
    n = int(max(cores, min(rate * args.cpu_seconds, batch.max_poas, len(base))))
    sample = base.slice(0, n)
    r = spoa_run(sample, cores)
    out = {"value": n / r.seconds, "unit": UNIT, "cores": cores, "kind": "reference",
           "gcups": r.cells / r.seconds / 1e9,
           "sample": "%d cfg2 windows, SPOA v3.0.0 (AVX2, full alignment) on %d threads, %.1f s" % (n, cores, r.seconds)}
    batch.reset()
    rc, added = batch.add_poa_groups_flat(sample.bases, sample.read_lengths, sample.window_n_reads)
    batch.generate_poa()
    cons, _, _ = batch.get_consensus()
    m = min(added, len(cons), n)
    if m > 0:
        out["consensus_agreement_with_engine"] = sum(1 for w in range(m) if cons[w] == r.consensus(w)) / m
        out["agreement_windows"] = m
    return out


def reference_cudapoa(args, base, device):
    """The reference cudapoa kernels (GenomeWorks 0.5.0, built unmodified for sm_100a) on the same GPU -- the kernel to
    beat -- driven as its own benchmark drives it: K concurrent batches on K streams from K host threads, staging under
    a mutex (cudapoa/benchmarks/multi_batch.hpp:43-56,157-170).  Reported only; outside every timed region of the
    native arm."""
    import oraclelib

    if not oraclelib.have_ref_cudapoa():
        return None
    import torch

    import parity

    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import reference_concurrent as rc

    lib = parity.ref_lib()
    ws = base.tile(args.ref_windows)
    shape = rc.SHAPES["cfg2"]
    best = None
    for k in (8, 16):
        free_b, _ = torch.cuda.mem_get_info(device)
        mem = int(free_b * 0.8 / k)
        dt, done, rounds, cap = rc.run_k(lib, ws, k, shape, mem)
        line = {"value": done / dt, "unit": UNIT, "k_concurrent_batches": k, "windows": done,
                "windows_per_batch": cap, "waves_per_batch": round(cap / (148 * 32.0), 2), "rounds": rounds}
        if best is None or line["value"] > best["value"]:
            best = line
    best["note"] = ("end to end through the same C ABI, best of K in {8, 16} concurrent batches (one stream and one host "
                    "thread each, staging under a mutex as in the reference's multi_batch benchmark); one-warp blocks: "
                    "a wave is 148 x 32 windows")
    return best


def bind_bench_hooks(dll):
    dll.cudapoa_b200_bench_relaunch.restype = ctypes.c_double
    dll.cudapoa_b200_bench_relaunch.argtypes = [ctypes.c_void_p, ctypes.c_int32]
    dll.cudapoa_b200_debug_dpx_peak.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_double),
                                                ctypes.POINTER(ctypes.c_double)]
    dll.cudapoa_b200_bench_relaunch_timed.restype = ctypes.c_double
    dll.cudapoa_b200_bench_relaunch_timed.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_double),
                                                      ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int32)]
    dll.cudapoa_b200_bench_phase_ms.restype = ctypes.c_int32
    dll.cudapoa_b200_bench_phase_ms.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.c_int32]


PHASES = ("fill", "traceback", "fusion", "topological_sort", "consensus")


def phase_ms(dll, batch, steps):
    """Per-phase milliseconds per step of the last relaunch_timed call (CUDA events between the launches)."""
    out = (ctypes.c_double * len(PHASES))()
    n = dll.cudapoa_b200_bench_phase_ms(batch._handle, out, len(PHASES))
    return {PHASES[i]: out[i] / steps for i in range(max(n, 0))}


def relaunch_timed(dll, batch, steps):
    fill_ms, graph_ms, n = ctypes.c_double(0), ctypes.c_double(0), ctypes.c_int32(0)
    total = dll.cudapoa_b200_bench_relaunch_timed(batch._handle, steps, ctypes.byref(fill_ms), ctypes.byref(graph_ms),
                                                  ctypes.byref(n))
    return total, fill_ms.value, graph_ms.value, n.value


def measure_other_config(name, spec, ws, lib, dll, device, ranks, max_mem, dpx_peak_gcups, world):
    """One of the other named shapes on this rank's GPU: first pass (stage, run, fetch: the gate), `steps` resident
    passes timed with CUDA events, `steps` end-to-end passes through one Batch with host buffers (wall clock)."""
    import torch

    from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

    batch = CudaPoaBatch(spec["max_reads"], spec["max_seq"], max_mem, output_type=spec["output"], device_id=device,
                         cuda_banded_alignment=spec.get("banded", True), alignment_band_width=256, lib=lib)
    try:
        if batch.max_poas < len(ws):
            ws = ws.slice(0, batch.max_poas)
        n = len(ws)

        def e2e_step():
            batch.reset()
            rc, added = batch.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
            if added != n:
                raise RuntimeError("%s: batch accepted %d of %d windows (status %d)" % (name, added, n, rc))
            batch.generate_poa()
            batch.fetch(spec["output"])

        e2e_step()
        statuses = batch.result_statuses()
        ok = sum(1 for s in statuses if s == 0)
        digest0 = batch.result_digest()
        stats = batch.stats()
        steps = spec["steps"]
        ranks.barrier()
        torch.cuda.synchronize()
        total_ms, fill_ms, graph_ms, launches = relaunch_timed(dll, batch, steps)
        phases = phase_ms(dll, batch, steps)
        ms_step = ranks.max(total_ms / steps)
        batch.fetch(spec["output"])
        stable = batch.result_digest() == digest0
        ranks.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        torch.cuda.synchronize()
        e2e_s = ranks.max((time.perf_counter() - t0) / steps)
        windows = ranks.sum(float(n))
        cells = ranks.sum(float(stats["dp_cells"]))
        gcups = cells / (ms_step * 1e-3) / 1e9
        fill_gcups = stats["dp_cells"] / (max(fill_ms / steps, 1e-9) * 1e-3) / 1e9
        return {"workload": spec["workload"], "batch_size": "BatchSize(%d, %d, 256)" % (spec["max_seq"], spec["max_reads"]),
                "windows_per_step_per_gpu": n, "steps": steps, "value": windows / (ms_step * 1e-3), "unit": UNIT,
                "ms_per_step": ms_step, "e2e": windows / e2e_s, "e2e_path": "one Batch: add_poa_group -> generate_poa -> get_*",
                "gcups": gcups, "fill_ms_per_step": fill_ms / steps, "graph_ms_per_step": graph_ms / steps,
                "phase_ms_per_step": phases, "launches_per_step": launches // max(steps, 1),
                "roofline_frac": (fill_gcups / dpx_peak_gcups) if dpx_peak_gcups > 0 and fill_ms > 0 else None,
                "whole_step_frac": (gcups / world / dpx_peak_gcups) if dpx_peak_gcups > 0 else None,
                "windows_ok": ok, "results_stable": stable, "exact_path_alignments": stats["exact_path"],
                "alignments": stats["reads_aligned"], "dp_cells_per_window": stats["dp_cells"] / max(n, 1)}
    finally:
        batch.close()


def run_native(args, rank, world, local_rank):
    # windows first: the generator forks a process pool, which must happen before CUDA is initialised
    ws, base = make_windows("cfg2", args.windows, args.distinct, rank * args.distinct)
    other_ws = {}
    if not args.no_other_configs:
        for name, spec in OTHER_CONFIGS.items():
            if spec["gen"] in ("cfg2", "cfg3"):
                other_ws[name] = ws.slice(0, spec["windows"])
            else:
                other_ws[name] = make_windows(spec["gen"], spec["windows"], args.distinct, rank * args.distinct)[0]

    import torch  # device sync + torch.distributed plumbing only

    from claragenomicsanalysis_b200 import _cabi, build
    from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the POA path has no CPU fallback")
    build.build()
    lib = _cabi.native()
    dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
    bind_bench_hooks(dll)

    device = local_rank
    torch.cuda.set_device(device)
    from claragenomicsanalysis_b200.distributed import Ranks

    ranks = Ranks(backend="nccl" if world > 1 else None)

    def barrier():
        ranks.barrier()
        torch.cuda.synchronize()

    max_over_ranks = ranks.max
    sum_over_ranks = ranks.sum

    free_b, total_b = torch.cuda.mem_get_info(device)
    max_mem = int(min(free_b * 0.85, args.max_mem_gb * (1 << 30)))
    batch = CudaPoaBatch(ENGINE["max_reads"], ENGINE["max_seq"], max_mem, output_type="consensus", device_id=device,
                         cuda_banded_alignment=True, alignment_band_width=ENGINE["band"], lib=lib)
    if batch.max_poas < len(ws):
        ws = ws.slice(0, batch.max_poas)
    n_win = len(ws)

    def e2e_step():
        batch.reset()
        rc, added = batch.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
        if added != n_win:
            raise RuntimeError("batch accepted %d of %d windows (status %d)" % (added, n_win, rc))
        batch.generate_poa()
        batch.fetch("consensus")

    sampler = ClockSampler(device)
    sampler.start()  # early: nvidia-smi needs seconds to start on a box with many GPUs; samples are windowed below
    # first pass: stage, run, fetch -- also the correctness gate of this run
    e2e_step()
    statuses = batch.result_statuses()
    n_ok = sum(1 for s in statuses if s == 0)
    digest0 = batch.result_digest()
    cons_chars = batch.result_total_chars()
    stats = batch.stats()
    if n_ok != n_win:
        raise RuntimeError("only %d of %d windows succeeded" % (n_ok, n_win))

    # ---- resident leg: inputs already in HBM, K passes timed with CUDA events on the batch's stream ----
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        dll.cudapoa_b200_bench_relaunch(batch._handle, 1)
    barrier()
    sampler.mark()
    ms_total, fill_total, graph_total, n_launches = relaunch_timed(dll, batch, args.steps)
    phases = phase_ms(dll, batch, args.steps)
    barrier()
    clocks = sampler.stop()
    if ms_total <= 0:
        raise RuntimeError("relaunch failed: " + lib.error())
    ms_step = max_over_ranks(ms_total / args.steps)
    launches_per_step = n_launches // args.steps
    fill_ms = fill_total / args.steps
    graph_ms = graph_total / args.steps
    batch.fetch("consensus")
    if batch.result_digest() != digest0:
        raise RuntimeError("results changed between launches")
    total_windows = sum_over_ranks(float(n_win))
    total_cells = sum_over_ranks(float(stats["dp_cells"]))
    value = total_windows / (ms_step * 1e-3)
    gcups = total_cells / (ms_step * 1e-3) / 1e9

    # ---- end-to-end leg: host buffers -> C ABI -> host results every step ----
    from claragenomicsanalysis_b200.pipeline import WindowPipeline

    batch.reset()
    chunk = min(args.chunk, n_win)
    pipe = WindowPipeline(ENGINE["max_reads"], ENGINE["max_seq"], int(args.worker_mem_gb * (1 << 30)),
                          devices=(device,), workers_per_device=args.workers, output_type="consensus",
                          cuda_banded_alignment=True, alignment_band_width=ENGINE["band"])
    for _ in range(warmup):
        pipe.run(ws, chunk)
    if pipe.digest() != digest0:
        raise RuntimeError("end-to-end results differ from the first pass")
    barrier()
    t0 = time.perf_counter()
    h2d = d2h = 0
    e2e_launches = 0
    for _ in range(args.steps):
        pipe.run(ws, chunk)
        st2 = pipe.stats()
        h2d += st2["h2d_bytes"]
        d2h += st2["d2h_bytes"]
        e2e_launches += st2["kernel_launches"]
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    e2e_s = max_over_ranks((t1 - t0) / args.steps)
    barrier()
    e2e_value = total_windows / e2e_s
    if pipe.digest() != digest0:
        raise RuntimeError("end-to-end results differ from the first pass")
    pipe.close()

    # ---- rooflines ----
    # Binding view: DP cells of the step / time of the dominant kernel (the fill launches, CUDA events
    # on the batch's stream) against the s16x2 DPX issue rate measured in this run (>= 1 such instruction per cell).
    # `traffic`: what ncu measured for the same launches (profiles/ncu_traffic.json), per launch; the algorithmic bytes
    # of a fill launch are the bases of the reads it aligns + their 4-byte lengths.
    hbm_peak, hbm_src = measured_peaks()
    alg_bytes = algorithmic_bytes(ws, cons_chars)
    n_fill = max(1, int(ws.window_n_reads.max()) - 1)
    first_reads = np.concatenate(([0], np.cumsum(ws.window_n_reads, dtype=np.int64)[:-1]))
    backbone_bytes = int(ws.read_lengths[first_reads].sum())
    fill_alg_bytes = (int(ws.n_bases) - backbone_bytes + 4 * (int(ws.n_reads) - n_win)) / n_fill  # per launch
    fill_launch_ms = fill_ms / n_fill
    achieved_gbs = fill_alg_bytes / (fill_launch_ms * 1e-3) / 1e9
    traffic = ncu_traffic()
    fill_traffic = traffic.get("poa_fill_kernel", {}).get("dram_bytes_per_window_step")
    graph_traffic = sum(traffic.get(k, {}).get("dram_bytes_per_window_step") or 0.0
                        for k in ("poa_trace_kernel", "poa_fuse_kernel", "poa_sort_kernel", "poa_final_kernel")) or None
    step_phase_ms = fill_ms + graph_ms
    r16 = ctypes.c_double(0)
    r16ms = ctypes.c_double(0)
    dll.cudapoa_b200_debug_dpx_peak(device, 0, ctypes.byref(r16), ctypes.byref(r16ms))
    dpx_peak_gcups = r16.value / 1e9  # >= 1 s16x2 DPX instruction per cell
    per_gpu_gcups = gcups / world
    fill_gcups = stats["dp_cells"] / (fill_ms * 1e-3) / 1e9
    traffic_per_launch = (fill_traffic * n_win / n_fill) if fill_traffic else None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "value_kind": "device-resident inputs, kernels only (CUDA events); e2e.value is the like-for-like number against "
                      "the reference arm and cpu_baseline",
        "dtype": "int16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "windows_per_step_per_gpu": n_win, "distinct_windows_per_gpu": len(base),
                   "batch_size": "BatchSize(640, 30, 256)", "l2": "inputs + per-window state larger than L2 (%.1f GB "
                   "touched per step)" % (n_win * 1.2e-3), "sharding": "windows split across ranks, no collective"},
        "gcups": gcups,
        "dp_cells_per_window": stats["dp_cells"] / n_win,
        "exact_path_alignments": stats["exact_path"], "alignments": stats["reads_aligned"],
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d // args.steps),
                "d2h_bytes_per_step": int(d2h // args.steps), "ms_per_step": e2e_s * 1e3,
                "path": "cudapoa_b200_pipeline_run(host buffers): %d workers x (add_poa_group -> generate_poa -> "
                        "get_consensus), chunks of %d windows" % (args.workers, chunk)},
        "gpu_launches": int(args.steps * launches_per_step + e2e_launches),
        "kernels": {
            "launches_per_step": int(launches_per_step),
            "poa_fill_kernel": {"launches": n_fill, "ms_per_step": fill_ms,
                                "share_of_step": fill_ms / step_phase_ms if step_phase_ms > 0 else None, "gcups": fill_gcups,
                                "dram_gb_per_s": (fill_traffic * n_win / (fill_ms * 1e-3) / 1e9) if fill_traffic else None,
                                "note": "includes the stand-by launches after it (poa_fill_rebased_kernel, "
                                        "poa_exact_kernel: nothing to do on this workload) and poa_init_kernel"},
            "graph_kernels": {"launches": max(int(launches_per_step) - 3 * n_fill - 1, 0), "ms_per_step": graph_ms,
                              "share_of_step": graph_ms / step_phase_ms if step_phase_ms > 0 else None,
                              "dram_gb_per_s": (graph_traffic * n_win / (graph_ms * 1e-3) / 1e9) if graph_traffic else None,
                              "poa_trace_kernel_ms": phases.get("traceback"), "poa_fuse_kernel_ms": phases.get("fusion"),
                              "poa_sort_kernel_ms": phases.get("topological_sort"),
                              "poa_final_kernel_ms": phases.get("consensus")},
            "note": "CUDA events between the launches of the timed region, this rank's GPU",
        },
        "roofline": {"bound": "dpx-int16", "kernel": "poa_fill_kernel", "achieved": fill_gcups, "peak": dpx_peak_gcups,
                     "unit": "GCUPS", "frac": fill_gcups / dpx_peak_gcups if dpx_peak_gcups > 0 else None,
                     "traffic": traffic_per_launch, "algorithmic_bytes_per_launch": fill_alg_bytes,
                     "launch_ms": fill_launch_ms, "whole_step_gcups": per_gpu_gcups,
                     "whole_step_frac": per_gpu_gcups / dpx_peak_gcups if dpx_peak_gcups > 0 else None,
                     "peak_source": "VIADDMNMX.S16x2 thread-instructions/s measured in this run "
                                    "(cudapoa_b200_debug_dpx_peak, %.2f ms): >= 1 per DP cell" % r16ms.value,
                     "traffic_source": "profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of this "
                                       "kernel's launches in the committed ncu capture, per launch, scaled to this "
                                       "step's windows"},
        "roofline_hbm": {"bound": "hbm", "kernel": "poa_fill_kernel", "achieved": achieved_gbs, "peak": hbm_peak,
                         "unit": "GB/s", "frac": achieved_gbs / hbm_peak, "traffic": traffic_per_launch,
                         "peak_source": hbm_src, "algorithmic_bytes_per_launch": fill_alg_bytes,
                         "algorithmic_bytes_per_step_all_kernels": alg_bytes,
                         "note": "secondary view: the path is not bound by its input bytes"},
        "clocks": clocks,
    }
    batch.close()
    batch = None

    if other_ws:
        line["configs"] = {}
        for name, spec in OTHER_CONFIGS.items():
            line["configs"][name] = guarded(measure_other_config, name, spec, other_ws[name], lib, dll, device, ranks,
                                            max_mem, dpx_peak_gcups, world)

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        def baseline_leg():
            b = CudaPoaBatch(ENGINE["max_reads"], ENGINE["max_seq"], min(max_mem, 16 << 30), output_type="consensus",
                             device_id=device, cuda_banded_alignment=True, alignment_band_width=ENGINE["band"], lib=lib)
            try:
                return cpu_baseline(args, base, b)
            finally:
                b.close()

        line["cpu_baseline"] = guarded(baseline_leg)
        with _StdoutToStderr():
            ref_gpu = guarded(reference_cudapoa, args, base, device)
        if ref_gpu is not None:
            line["reference_cudapoa_gpu"] = ref_gpu
    if rank == 0:
        print(json.dumps(line))
    ranks.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--windows", type=int, default=18944, help="windows per step per GPU (8 waves of 148 x 16)")
    ap.add_argument("--chunk", type=int, default=1184, help="end-to-end leg: windows per batch round")
    ap.add_argument("--workers", type=int, default=8, help="end-to-end leg: host threads / streams / batches")
    ap.add_argument("--worker-mem-gb", type=float, default=6.0)
    ap.add_argument("--distinct", type=int, default=65536, help="distinct generated windows per shape (SURVEY.md 8(d)); "
                    "tiled when a step has more")
    ap.add_argument("--max-mem-gb", type=float, default=64.0)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--ref-windows", type=int, default=16384, help="windows of the reference cudapoa leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    args = ap.parse_args()
    rank = env_int("RANK", 0)
    world = env_int("WORLD_SIZE", 1)
    local_rank = env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_native(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
