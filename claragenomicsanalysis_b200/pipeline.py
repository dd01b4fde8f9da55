"""Window pipeline: many windows through several batches/streams (and devices) of the native engine in one call.

Python face of the ``cudapoa_b200_pipeline_*`` entry points (include/cudapoa_b200.h): the stream scheduler and
sharding layer above the cudapoa Batch interface.  Results are indexed by window, whatever worker processed them."""
import ctypes

import numpy as np

from . import _cabi
from ._cabi import OUTPUT_CONSENSUS, OUTPUT_MSA, BatchSizeC, StatsC
from .cudapoa import band_mode_value

_configured = set()


def _dll(lib=None):
    lib = lib or _cabi.native()
    dll = lib._dll
    if id(dll) not in _configured:
        dll.cudapoa_b200_pipeline_last_error.restype = ctypes.c_char_p
        dll.cudapoa_b200_pipeline_create.restype = ctypes.c_void_p
        dll.cudapoa_b200_pipeline_create.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_size_t,
                                                     ctypes.c_int8, ctypes.POINTER(BatchSizeC), ctypes.c_int16,
                                                     ctypes.c_int16, ctypes.c_int16, ctypes.c_int32]
        dll.cudapoa_b200_pipeline_destroy.argtypes = [ctypes.c_void_p]
        dll.cudapoa_b200_pipeline_run.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int32, ctypes.c_int32]
        for name in ("workers", "capacity"):
            getattr(dll, "cudapoa_b200_pipeline_" + name).argtypes = [ctypes.c_void_p]
        dll.cudapoa_b200_pipeline_wall_ms.argtypes = [ctypes.c_void_p]
        dll.cudapoa_b200_pipeline_wall_ms.restype = ctypes.c_double
        dll.cudapoa_b200_pipeline_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(StatsC)]
        dll.cudapoa_b200_pipeline_status.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        dll.cudapoa_b200_pipeline_consensus.argtypes = [ctypes.c_void_p, ctypes.c_int32,
                                                        ctypes.POINTER(ctypes.c_void_p),
                                                        ctypes.POINTER(ctypes.c_void_p)]
        dll.cudapoa_b200_pipeline_msa_rows.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        dll.cudapoa_b200_pipeline_msa_row.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                                      ctypes.POINTER(ctypes.c_void_p)]
        dll.cudapoa_b200_pipeline_total_chars.argtypes = [ctypes.c_void_p]
        dll.cudapoa_b200_pipeline_total_chars.restype = ctypes.c_int64
        dll.cudapoa_b200_pipeline_digest.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32]
        dll.cudapoa_b200_pipeline_digest.restype = ctypes.c_uint64
        _configured.add(id(dll))
    return dll


def shard_bounds(n_items, n_shards, shard):
    """Contiguous, balanced split of ``n_items`` over ``n_shards``: bounds [lo, hi) of shard ``shard``.
    Used to give every rank (one process per GPU) its own windows; no rank needs anything from another."""
    base, rem = divmod(int(n_items), int(n_shards))
    lo = shard * base + min(shard, rem)
    return lo, lo + base + (1 if shard < rem else 0)


class WindowPipeline:
    def __init__(self, max_sequences_per_poa, max_sequence_size, mem_per_worker, devices=(0,), workers_per_device=2,
                 output_type="consensus", gap_score=-8, mismatch_score=-6, match_score=8, cuda_banded_alignment=True,
                 alignment_band_width=256, lib=None):
        lib = lib or _cabi.native()
        self._dll = _dll(lib)
        self._mask = OUTPUT_CONSENSUS if output_type == "consensus" else OUTPUT_MSA
        bs = BatchSizeC()
        lib.check(lib.batch_size_default(ctypes.byref(bs), max_sequence_size, max_sequences_per_poa,
                                         alignment_band_width), "BatchSize")
        self.batch_size = bs
        devs = np.ascontiguousarray(list(devices), dtype=np.int32)
        h = self._dll.cudapoa_b200_pipeline_create(devs.ctypes.data, int(devs.shape[0]), workers_per_device,
                                                   int(mem_per_worker), self._mask, ctypes.byref(bs), gap_score,
                                                   mismatch_score, match_score, band_mode_value(cuda_banded_alignment))
        if not h:
            raise RuntimeError(self._dll.cudapoa_b200_pipeline_last_error().decode("utf-8", "replace"))
        self._h = ctypes.c_void_p(h)
        self.n_windows = 0

    def close(self):
        if getattr(self, "_h", None):
            self._dll.cudapoa_b200_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def workers(self):
        return self._dll.cudapoa_b200_pipeline_workers(self._h)

    @property
    def capacity(self):
        return self._dll.cudapoa_b200_pipeline_capacity(self._h)

    def run(self, ws, chunk_windows=0):
        """All windows of a WindowSet through the workers (blocking).  Host buffers in, host results out."""
        w_ptr = ws.weights.ctypes.data if ws.weights is not None else None
        rc = self._dll.cudapoa_b200_pipeline_run(self._h, ws.bases.ctypes.data, w_ptr, ws.read_lengths.ctypes.data,
                                                 ws.window_n_reads.ctypes.data, len(ws), chunk_windows)
        if rc != 0:
            raise RuntimeError(self._dll.cudapoa_b200_pipeline_last_error().decode("utf-8", "replace"))
        self.n_windows = len(ws)

    @property
    def wall_ms(self):
        return self._dll.cudapoa_b200_pipeline_wall_ms(self._h)

    def stats(self):
        out = StatsC()
        self._dll.cudapoa_b200_pipeline_stats(self._h, ctypes.byref(out))
        return {name: getattr(out, name) for name, _ in StatsC._fields_}

    def statuses(self):
        return [self._dll.cudapoa_b200_pipeline_status(self._h, i) for i in range(self.n_windows)]

    def consensus(self, i):
        seq, cov = ctypes.c_void_p(), ctypes.c_void_p()
        n = self._dll.cudapoa_b200_pipeline_consensus(self._h, i, ctypes.byref(seq), ctypes.byref(cov))
        if n <= 0:
            return "", []
        return (ctypes.string_at(seq.value, n).decode("ascii"), list((ctypes.c_uint16 * n).from_address(cov.value)))

    def msa(self, i):
        rows = []
        seq = ctypes.c_void_p()
        for r in range(self._dll.cudapoa_b200_pipeline_msa_rows(self._h, i)):
            n = self._dll.cudapoa_b200_pipeline_msa_row(self._h, i, r, ctypes.byref(seq))
            rows.append(ctypes.string_at(seq.value, n).decode("ascii"))
        return rows

    def total_chars(self):
        return int(self._dll.cudapoa_b200_pipeline_total_chars(self._h))

    def digest(self, first=0, count=None):
        return int(self._dll.cudapoa_b200_pipeline_digest(self._h, first, self.n_windows if count is None else count))


def assembly(consensus, coverage):
    """Concatenate window consensuses into one sequence, trimming each window at both ends while the per-base
    coverage is below the window's average coverage -- the reference's end-to-end check
    (cudapoa/benchmarks/multi_batch.hpp:176-207; integer average, a window whose trimmed range is empty or a single
    base contributes nothing, exactly as there)."""
    genome = []
    for cons, cov in zip(consensus, coverage):
        n = len(cov)
        if n == 0:
            continue
        average = int(sum(int(c) for c in cov)) // n
        begin, end = 0, len(cons) - 1
        while begin < n and cov[begin] < average:
            begin += 1
        while end >= 0 and cov[end] < average:
            end -= 1
        if begin < end:
            genome.append(cons[begin:end + 1])
    return "".join(genome)
