#!/usr/bin/env python
"""Benchmark of the B200 cudapoa engine on BASELINE.json's metric: POA windows/s (+ GCUPS).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

Workload at every N: BASELINE.json configs[1] -- Racon-style polishing windows, 500 bp template x 30 ONT-like reads,
static band 256, int16 scores, consensus output (SURVEY.md 8(d) "cfg 2").  One "step" = one batch of `--windows`
such windows per GPU through the hot path.  Windows are independent: each rank owns its own windows (weak scaling),
no collective on the data path; torch.distributed is used for the barrier and the max-over-ranks only.

Native arm, three numbers per run:
  value      windows/s with the batch's inputs already resident in HBM: K launches of the POA kernel timed with CUDA
             events on the batch's stream (cudapoa_b200_bench_relaunch), max over ranks.
  e2e        windows/s through the C ABI with HOST buffers every step: the window pipeline (cudapoa_b200_pipeline_run)
             drives `--workers` Batch objects, each reset -> add_poa_group (host staging) -> generate_poa (H2D + kernel)
             -> get_consensus (D2H + decode) per chunk of windows; wall clock between device syncs.
  roofline   HBM view (algorithmic bytes of SURVEY.md 8(d) / kernel time vs MEASURED_PEAKS.json) and, in
             `roofline_dpx`, the binding DPX view: achieved GCUPS vs the s16x2 DPX issue rate measured in this run.
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


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


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
        t_begin = getattr(self, "t_begin", 0.0)
        rows = [r for r, t in zip(self.rows, self.stamps) if t_begin <= t <= t_end + 0.05]
        window = "timed region"
        if not rows:  # region shorter than the sampling period: the warm-up steps just before ran the same load
            rows = [r for r, t in zip(self.rows, self.stamps) if t >= t_begin - 2.0]
            window = "last 2 s of warm-up + timed region (same load; no sample fell inside the timed region)"
        sm = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in rows if r[3].replace(".", "").isdigit()]
        reasons = set()
        self.window = window
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "window": self.window}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def make_windows(n_windows, distinct, start_index):
    from claragenomicsanalysis_b200.windows import config_windows

    base = config_windows("cfg2", min(distinct, n_windows), start_index=start_index)
    return base.tile(n_windows), base


def algorithmic_bytes(ws, total_consensus_chars):
    """bases + 32 B descriptor per window + 2 B per read length + consensus (1 B) and coverage (2 B)
    per consensus base.  Weights are not transferred when every weight is 1."""
    return int(ws.n_bases + 32 * len(ws) + 2 * ws.n_reads + 3 * total_consensus_chars)


def spoa_run(ws, threads=0):
    import oraclelib

    return oraclelib.SpoaRun(ws, match=8, mismatch=-6, gap=-8, msa=False, threads=threads)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """SPOA (the reference repository's CPU POA) on all host cores; rank 0 only."""
    if rank != 0:
        return
    import oraclelib

    if not oraclelib.have_ref_spoa():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libspoa_ref.so was not built"}))
        return
    cores = host_cores()
    probe, _ = make_windows(max(2 * cores, 8), 4096, 0)
    r = spoa_run(probe, cores)
    rate = len(probe) / max(r.seconds, 1e-6)
    budget = min(10.0, 150.0 / max(args.steps + args.warmup, 1))
    n = max(cores, int(rate * budget))
    ws, _ = make_windows(n, 4096, 0)
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
        "config": {"workload": WORKLOAD, "windows_per_step": n, "engine": "SPOA v3.0.0 kNW linear gaps, full (unbanded)"},
        "gcups": cells / t / 1e9,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": "%d cfg2 windows per step x %d steps, SPOA on %d threads" % (n, args.steps, cores)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def ncu_traffic():
    """DRAM bytes per window and step of each kernel, from the committed ncu launch list (profiles/ncu_traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum summed over the launches of one step, divided by its windows)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def run_native(args, rank, world, local_rank):
    import torch  # device sync + torch.distributed plumbing only

    from claragenomicsanalysis_b200 import _cabi, build
    from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the POA path has no CPU fallback")
    build.build()
    lib = _cabi.native()
    dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
    dll.cudapoa_b200_bench_relaunch.restype = ctypes.c_double
    dll.cudapoa_b200_bench_relaunch.argtypes = [ctypes.c_void_p, ctypes.c_int32]
    dll.cudapoa_b200_debug_dpx_peak.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_double),
                                                ctypes.POINTER(ctypes.c_double)]
    dll.cudapoa_b200_bench_relaunch_timed.restype = ctypes.c_double
    dll.cudapoa_b200_bench_relaunch_timed.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_double),
                                                      ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int32)]

    device = local_rank
    torch.cuda.set_device(device)
    from claragenomicsanalysis_b200.distributed import Ranks

    ranks = Ranks(backend="nccl" if world > 1 else None)

    def barrier():
        ranks.barrier()
        torch.cuda.synchronize()

    max_over_ranks = ranks.max
    sum_over_ranks = ranks.sum

    ws, base = make_windows(args.windows, args.distinct, rank * args.distinct)
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

    # ---- resident leg: inputs already in HBM, K launches timed with CUDA events on the batch's stream ----
    for _ in range(max(args.warmup, 3)):
        dll.cudapoa_b200_bench_relaunch(batch._handle, 1)
    barrier()
    sampler.mark()
    # the timed region: K steps, CUDA events on the batch's stream around the whole region and between its launches
    fill_ms = ctypes.c_double(0)
    graph_ms = ctypes.c_double(0)
    n_launches = ctypes.c_int32(0)
    ms_total = dll.cudapoa_b200_bench_relaunch_timed(batch._handle, args.steps, ctypes.byref(fill_ms),
                                                     ctypes.byref(graph_ms), ctypes.byref(n_launches))
    barrier()
    clocks = sampler.stop()
    if ms_total <= 0:
        raise RuntimeError("relaunch failed: " + lib.error())
    ms_step = max_over_ranks(ms_total / args.steps)
    # per-kernel times of the timed region, per step
    launches_per_step = n_launches.value // args.steps
    fill_ms = ctypes.c_double(fill_ms.value / args.steps)
    graph_ms = ctypes.c_double(graph_ms.value / args.steps)
    batch.fetch("consensus")
    if batch.result_digest() != digest0:
        raise RuntimeError("results changed between launches")
    total_windows = sum_over_ranks(float(n_win))
    total_cells = sum_over_ranks(float(stats["dp_cells"]))
    value = total_windows / (ms_step * 1e-3)
    gcups = total_cells / (ms_step * 1e-3) / 1e9

    # ---- end-to-end leg: host buffers -> C ABI -> host results every step ----
    # The window pipeline (stream scheduler above the Batch interface): `--workers` host threads, each with its own
    # stream and Batch, stage / launch / collect chunks of one wave of windows; staging overlaps other chunks' kernels.
    from claragenomicsanalysis_b200.pipeline import WindowPipeline

    batch.reset()
    chunk = min(args.chunk, n_win)
    pipe = WindowPipeline(ENGINE["max_reads"], ENGINE["max_seq"], int(args.worker_mem_gb * (1 << 30)),
                          devices=(device,), workers_per_device=args.workers, output_type="consensus",
                          cuda_banded_alignment=True, alignment_band_width=ENGINE["band"])
    for _ in range(max(args.warmup, 3)):
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
    # The dominant kernel is poa_fill_kernel (one launch per read; share of the step measured above).  HBM view:
    # algorithmic bytes of a fill launch = the bases of the reads it aligns + their 4-byte lengths (SURVEY.md 8(d)'s
    # per-window figure restricted to what this kernel consumes; everything else it touches -- graph, descriptors,
    # score rows -- is intermediate state), against the launch durations measured with CUDA events.  `traffic` is what
    # ncu measured for the same launches: the streamed score rows.
    hbm_peak, hbm_src = measured_peaks()
    alg_bytes = algorithmic_bytes(ws, cons_chars)
    n_fill = max(1, int(ws.window_n_reads.max()) - 1)
    first_reads = np.concatenate(([0], np.cumsum(ws.window_n_reads, dtype=np.int64)[:-1]))
    backbone_bytes = int(ws.read_lengths[first_reads].sum())
    fill_alg_bytes = (int(ws.n_bases) - backbone_bytes + 4 * (int(ws.n_reads) - n_win)) / n_fill  # per launch
    fill_launch_ms = fill_ms.value / n_fill
    achieved_gbs = fill_alg_bytes / (fill_launch_ms * 1e-3) / 1e9
    traffic = ncu_traffic()
    fill_traffic = traffic.get("poa_fill_kernel", {}).get("dram_bytes_per_window_step")
    graph_traffic = traffic.get("poa_graph_kernel", {}).get("dram_bytes_per_window_step")
    phase_ms = fill_ms.value + graph_ms.value
    r16 = ctypes.c_double(0)
    r16ms = ctypes.c_double(0)
    dll.cudapoa_b200_debug_dpx_peak(device, 0, ctypes.byref(r16), ctypes.byref(r16ms))
    dpx_peak_gcups = r16.value / 1e9  # >= 1 s16x2 DPX instruction per cell
    per_gpu_gcups = gcups / world

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "windows_per_step_per_gpu": n_win, "distinct_windows_per_gpu": len(base),
                   "batch_size": "BatchSize(640, 30, 256)", "l2": "inputs + per-window state larger than L2 (%.1f GB "
                   "touched per step)" % (n_win * 2.7e-3), "sharding": "windows split across ranks, no collective"},
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
            "poa_fill_kernel": {"launches": n_fill, "ms_per_step": fill_ms.value,
                                "share_of_step": fill_ms.value / phase_ms if phase_ms > 0 else None,
                                "gcups": stats["dp_cells"] / (fill_ms.value * 1e-3) / 1e9,
                                "dram_gb_per_s": (fill_traffic * n_win / (fill_ms.value * 1e-3) / 1e9) if fill_traffic else None},
            "poa_fill_rebased_kernel + poa_exact_kernel": {
                "launches": 2 * n_fill if launches_per_step > 1 else 0,
                "note": "after every fill launch: long windows (rebased rows) and reads the packed fill declines; "
                        "neither occurs on this workload, their time is counted with the fill launches"},
            "poa_init_kernel": {"launches": 1 if launches_per_step > 1 else 0,
                                "note": "state + backbone, first launch of a step; counted with the fill launches"},
            "poa_graph_kernel": {"launches": max(int(launches_per_step) - 3 * n_fill - 1, 0), "ms_per_step": graph_ms.value,
                                 "share_of_step": graph_ms.value / phase_ms if phase_ms > 0 else None,
                                 "dram_gb_per_s": (graph_traffic * n_win / (graph_ms.value * 1e-3) / 1e9) if graph_traffic else None},
            "note": "CUDA events between the launches of the timed region, this rank's GPU",
        },
        "roofline": {"bound": "hbm", "kernel": "poa_fill_kernel", "achieved": achieved_gbs, "peak": hbm_peak,
                     "unit": "GB/s", "frac": achieved_gbs / hbm_peak,
                     "traffic": (fill_traffic * n_win / n_fill) if fill_traffic else None, "peak_source": hbm_src,
                     "traffic_source": "profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of this "
                                       "kernel's launches in the committed ncu capture, per launch, scaled to this "
                                       "step's windows",
                     "algorithmic_bytes_per_launch": fill_alg_bytes, "launch_ms": fill_launch_ms,
                     "algorithmic_bytes_per_step_all_kernels": alg_bytes,
                     "note": "integer DP: the binding roofline is the DPX issue rate, see roofline_dpx; the DRAM "
                             "traffic is the score matrix streamed for the traceback, not input"},
        "roofline_dpx": {"bound": "dpx-int16", "kernel": "poa_fill_kernel",
                         "achieved": stats["dp_cells"] / (fill_ms.value * 1e-3) / 1e9, "peak": dpx_peak_gcups,
                         "unit": "GCUPS",
                         "frac": stats["dp_cells"] / (fill_ms.value * 1e-3) / 1e9 / dpx_peak_gcups if dpx_peak_gcups > 0 else None,
                         "whole_step_gcups": per_gpu_gcups,
                         "whole_step_frac": per_gpu_gcups / dpx_peak_gcups if dpx_peak_gcups > 0 else None,
                         "peak_source": "VIADDMNMX.S16x2 thread-instructions/s measured in this run "
                                        "(cudapoa_b200_debug_dpx_peak, %.2f ms)" % r16ms.value},
        "clocks": clocks,
    }

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args, base, batch, lib)
        with _StdoutToStderr():
            ref_gpu = reference_cudapoa(args, base)
        if ref_gpu is not None:
            line["reference_cudapoa_gpu"] = ref_gpu
    batch.close()
    if rank == 0:
        print(json.dumps(line))
    ranks.close()


def cpu_baseline(args, base, batch, lib):
    """SPOA on all host cores over a bounded sample of the same windows + consensus agreement with the engine."""
    import oraclelib

    if not oraclelib.have_ref_spoa():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "libspoa_ref.so not built"}
    cores = host_cores()
    probe = base.slice(0, max(2 * cores, 8))
    r = spoa_run(probe, cores)
    rate = len(probe) / max(r.seconds, 1e-6)
    n = int(max(cores, rate * args.cpu_seconds))
    sample = base.tile(n)
    r = spoa_run(sample, cores)
    # agreement: the engine's consensus for the same windows
    batch.reset()
    batch.add_poa_groups_flat(sample.bases, sample.read_lengths, sample.window_n_reads)
    batch.generate_poa()
    cons, _, _ = batch.get_consensus()
    agree = sum(1 for w in range(n) if cons[w] == r.consensus(w))
    return {"value": n / r.seconds, "unit": UNIT, "cores": cores, "kind": "reference",
            "gcups": r.cells / r.seconds / 1e9,
            "sample": "%d cfg2 windows, SPOA v3.0.0 (AVX2, full alignment) on %d threads, %.1f s" % (n, cores, r.seconds),
            "consensus_agreement_with_engine": agree / n}


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


def reference_cudapoa(args, base):
    """The reference cudapoa kernels (GenomeWorks 0.5.0, built unmodified for sm_100a) on the same GPU: the kernel
    to beat.  Reported only; not part of any timed region of the native arm."""
    import oraclelib

    if not oraclelib.have_ref_cudapoa():
        return None
    try:
        import parity

        from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

        n = 4736
        ws = base.tile(n)
        b = CudaPoaBatch(ENGINE["max_reads"], ENGINE["max_seq"], 8 << 30, output_type="consensus",
                         cuda_banded_alignment=True, alignment_band_width=ENGINE["band"], lib=parity.ref_lib())
        best = None
        for _ in range(3):
            b.reset()
            t0 = time.perf_counter()
            rc, added = b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
            b.generate_poa()
            b.fetch("consensus")
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        b.close()
        return {"value": added / best, "unit": UNIT, "windows": added,
                "note": "end-to-end through the same C ABI, best of 3, one batch on one stream"}
    except Exception as e:  # the comparison is optional
        return {"error": str(e)[:200]}


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
    ap.add_argument("--distinct", type=int, default=1024, help="distinct generated windows (tiled to --windows)")
    ap.add_argument("--max-mem-gb", type=float, default=64.0)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
