"""ctypes access to the checkers under oracle/ (CPU restatement, reference SPOA, reference cudapoa build).
Test infrastructure: imported only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libcudapoa_oracle.so")
REF_CUDAPOA_SO = os.path.join(ORACLE_DIR, "_ref", "libcudapoa_ref.so")
REF_SPOA_SO = os.path.join(ORACLE_DIR, "_ref", "libspoa_ref.so")
RCP_TABLE = os.path.join(ROOT, "tests", "golden", "b200_mufu_rcp_u16.npy")


class OracleParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "max_sequence_size", "max_consensus_size", "max_nodes_per_window", "max_nodes_per_window_banded",
        "max_matrix_graph_dimension", "max_matrix_graph_dimension_banded", "max_matrix_sequence_dimension",
        "alignment_band_width", "max_sequences_per_poa", "gap_score", "mismatch_score", "match_score", "banded",
        "output_mask", "spoa_accurate", "stale_fill")] + [("rcp_table", ctypes.c_void_p),
                                                           ("rcp_table_len", ctypes.c_int32),
                                                           ("adaptive_band", ctypes.c_int32)]


def build_oracle():
    src = os.path.join(ORACLE_DIR, "cudapoa_oracle.cpp")
    if (not os.path.exists(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


_oracle = None
_rcp = None


def rcp_table():
    """MUFU.RCP table measured on a B200 (tests/golden/make_rcp_table.py), or None if not generated yet."""
    global _rcp
    if _rcp is None and os.path.exists(RCP_TABLE):
        _rcp = np.ascontiguousarray(np.load(RCP_TABLE), dtype=np.float32)
    return _rcp


def oracle():
    global _oracle
    if _oracle is None:
        lib = ctypes.CDLL(build_oracle())
        lib.oracle_run.restype = ctypes.c_void_p
        lib.oracle_run.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int32, ctypes.POINTER(OracleParams)]
        lib.oracle_free.argtypes = [ctypes.c_void_p]
        for name in ("oracle_score_bits", "oracle_size_bits"):
            getattr(lib, name).argtypes = [ctypes.c_void_p]
        for name in ("oracle_dp_cells", "oracle_graph_rows", "oracle_rcp_misses"):
            getattr(lib, name).argtypes = [ctypes.c_void_p]
            getattr(lib, name).restype = ctypes.c_int64
        lib.oracle_status.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        lib.oracle_consensus.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p),
                                         ctypes.POINTER(ctypes.c_void_p)]
        lib.oracle_msa_rows.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        lib.oracle_msa_row.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                       ctypes.POINTER(ctypes.c_void_p)]
        lib.oracle_graph_nodes.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32]
        lib.oracle_graph_edges.argtypes = [ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 3 + [ctypes.c_int32]
        lib.oracle_default_params.argtypes = [ctypes.POINTER(OracleParams)] + [ctypes.c_int32] * 3
        _oracle = lib
    return _oracle


def make_params(max_seq=1024, max_reads=100, band=256, banded=True, output="consensus", gap=-8, mismatch=-6,
                match=8, spoa_accurate=False, stale_fill=0, use_rcp=True, batch_size=None, adaptive=False):
    p = OracleParams()
    oracle().oracle_default_params(ctypes.byref(p), max_seq, max_reads, band)
    if batch_size is not None:  # a BatchSizeC or dict with the nine BatchSize fields
        d = batch_size if isinstance(batch_size, dict) else batch_size.as_dict()
        for k, v in d.items():
            setattr(p, k, v)
    p.gap_score, p.mismatch_score, p.match_score = gap, mismatch, match
    p.banded = 1 if banded else 0
    p.output_mask = {"consensus": 1, "msa": 2, "both": 3}[output]
    p.spoa_accurate = 1 if spoa_accurate else 0
    p.stale_fill = stale_fill
    p.adaptive_band = 1 if adaptive else 0
    t = rcp_table() if use_rcp else None
    if t is not None:
        p.rcp_table = t.ctypes.data
        p.rcp_table_len = int(t.shape[0])
    return p


class OracleRun:
    """Results of oracle_run over a WindowSet."""

    def __init__(self, ws, params):
        self._lib = oracle()
        self._params = params
        w_ptr = ws.weights.ctypes.data if ws.weights is not None else None
        self._h = ctypes.c_void_p(self._lib.oracle_run(ws.bases.ctypes.data, w_ptr, ws.read_lengths.ctypes.data,
                                                       ws.window_n_reads.ctypes.data, len(ws), ctypes.byref(params)))
        self.n = len(ws)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.oracle_free(self._h)
            self._h = None

    @property
    def score_bits(self):
        return self._lib.oracle_score_bits(self._h)

    @property
    def size_bits(self):
        return self._lib.oracle_size_bits(self._h)

    @property
    def dp_cells(self):
        return self._lib.oracle_dp_cells(self._h)

    @property
    def graph_rows(self):
        return self._lib.oracle_graph_rows(self._h)

    @property
    def rcp_misses(self):
        return self._lib.oracle_rcp_misses(self._h)

    @property
    def widened(self):
        return self._lib.oracle_widened(self._h)

    def status(self, w):
        return self._lib.oracle_status(self._h, w)

    def statuses(self):
        return [self.status(w) for w in range(self.n)]

    def consensus(self, w):
        seq, cov = ctypes.c_void_p(), ctypes.c_void_p()
        n = self._lib.oracle_consensus(self._h, w, ctypes.byref(seq), ctypes.byref(cov))
        if n <= 0:
            return "", []
        return (ctypes.string_at(seq.value, n).decode("ascii"), list((ctypes.c_uint16 * n).from_address(cov.value)))

    def msa(self, w):
        rows = []
        seq = ctypes.c_void_p()
        for r in range(self._lib.oracle_msa_rows(self._h, w)):
            n = self._lib.oracle_msa_row(self._h, w, r, ctypes.byref(seq))
            rows.append(ctypes.string_at(seq.value, n).decode("ascii"))
        return rows

    def graph(self, w):
        n = self._lib.oracle_graph_nodes(self._h, w, None, 0)
        labels = ctypes.create_string_buffer(max(n, 1))
        self._lib.oracle_graph_nodes(self._h, w, labels, n)
        m = self._lib.oracle_graph_edges(self._h, w, None, None, None, 0)
        src, dst, wgt = (np.zeros(max(m, 1), np.int32) for _ in range(3))
        self._lib.oracle_graph_edges(self._h, w, src.ctypes.data, dst.ctypes.data, wgt.ctypes.data, m)
        return labels.raw[:n], [(int(src[k]), int(dst[k]), int(wgt[k])) for k in range(m)]


def have_ref_spoa():
    return os.path.exists(REF_SPOA_SO)


def have_ref_cudapoa():
    return os.path.exists(REF_CUDAPOA_SO)


_spoa = None


def spoa():
    global _spoa
    if _spoa is None:
        lib = ctypes.CDLL(REF_SPOA_SO)
        lib.spoa_ref_run.restype = ctypes.c_void_p
        lib.spoa_ref_run.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int32] * 6
        lib.spoa_ref_free.argtypes = [ctypes.c_void_p]
        lib.spoa_ref_seconds.argtypes = [ctypes.c_void_p]
        lib.spoa_ref_seconds.restype = ctypes.c_double
        lib.spoa_ref_cells.argtypes = [ctypes.c_void_p]
        lib.spoa_ref_cells.restype = ctypes.c_int64
        lib.spoa_ref_threads.argtypes = [ctypes.c_void_p]
        lib.spoa_ref_consensus.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p)]
        lib.spoa_ref_msa_rows.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        lib.spoa_ref_msa_row.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.POINTER(ctypes.c_void_p)]
        _spoa = lib
    return _spoa


class SpoaRun:
    """Reference SPOA (kNW, linear gaps) over a WindowSet on `threads` host threads (0 = all)."""

    def __init__(self, ws, match=8, mismatch=-6, gap=-8, msa=False, threads=0):
        self._lib = spoa()
        self._h = ctypes.c_void_p(self._lib.spoa_ref_run(ws.bases.ctypes.data, ws.read_lengths.ctypes.data,
                                                         ws.window_n_reads.ctypes.data, len(ws), match, mismatch,
                                                         gap, 1 if msa else 0, threads))
        self.n = len(ws)
        self.seconds = self._lib.spoa_ref_seconds(self._h)
        self.cells = self._lib.spoa_ref_cells(self._h)
        self.threads = self._lib.spoa_ref_threads(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.spoa_ref_free(self._h)
            self._h = None

    def consensus(self, w):
        seq = ctypes.c_void_p()
        n = self._lib.spoa_ref_consensus(self._h, w, ctypes.byref(seq))
        return ctypes.string_at(seq.value, n).decode("ascii") if n > 0 else ""

    def msa(self, w):
        rows = []
        seq = ctypes.c_void_p()
        for r in range(self._lib.spoa_ref_msa_rows(self._h, w)):
            n = self._lib.spoa_ref_msa_row(self._h, w, r, ctypes.byref(seq))
            rows.append(ctypes.string_at(seq.value, n).decode("ascii"))
        return rows
