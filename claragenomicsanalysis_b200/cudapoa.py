"""Python face of the B200 cudapoa engine: same class, arguments and return values as pygenomeworks'
``genomeworks.cudapoa.CudaPoaBatch`` (pygenomeworks/genomeworks/cudapoa/cudapoa.pyx:60-298), implemented over the
C ABI of include/cudapoa_b200.h instead of Cython.  Flat-buffer helpers (``add_poa_groups_flat``) are additions for
bulk feeding from numpy arrays.
"""
import ctypes

import numpy as np

from . import _cabi
from ._cabi import OUTPUT_CONSENSUS, OUTPUT_MSA, STATUS_NAMES, BatchSizeC, StatsC

# StatusType values, importable like `from genomeworks.cudapoa import ...` constants in the Cython module
(success, exceeded_maximum_poas, exceeded_maximum_sequence_size, exceeded_maximum_sequences_per_poa,
 node_count_exceeded_maximum_graph_size, edge_count_exceeded_maximum_graph_size,
 seq_len_exceeded_maximum_nodes_per_window, loop_count_exceeded_upper_bound, output_type_unavailable,
 generic_error) = range(10)


def status_to_str(status):
    """Convert status to its string representation (cudapoa.pyx:33-57)."""
    if 0 <= int(status) < len(STATUS_NAMES):
        return STATUS_NAMES[int(status)]
    raise RuntimeError("Unknown error status : " + str(status))


def init(lib=None):
    """cudapoa::Init() (cudapoa.hpp:41)."""
    lib = lib or _cabi.native()
    return lib.check(lib.init(), "Init")


def batch_size(max_sequence_size=1024, max_sequences_per_poa=100, alignment_band_width=256, lib=None):
    """BatchSize three-argument constructor (batch.hpp:75-96)."""
    lib = lib or _cabi.native()
    out = BatchSizeC()
    lib.check(lib.batch_size_default(ctypes.byref(out), max_sequence_size, max_sequences_per_poa,
                                     alignment_band_width), "BatchSize")
    return out


def band_mode_value(cuda_banded_alignment):
    """False/True as in the reference binding; "adaptive" (or 2) selects this engine's adaptive banding."""
    if cuda_banded_alignment == "adaptive" or cuda_banded_alignment == 2:
        return 2
    return 1 if cuda_banded_alignment else 0


class CudaPoaBatch:
    """Python API for CUDA-accelerated partial order alignment (mirror of pygenomeworks' CudaPoaBatch)."""

    def __init__(self,
                 max_sequences_per_poa,
                 max_sequence_size,
                 max_gpu_mem,
                 output_type="consensus",
                 device_id=0,
                 stream=None,
                 gap_score=-8,
                 mismatch_score=-6,
                 match_score=8,
                 cuda_banded_alignment=False,
                 alignment_band_width=256,
                 max_consensus_size=None,
                 max_nodes_per_window=None,
                 max_nodes_per_window_banded=None,
                 batch_size_c=None,
                 lib=None):
        self._lib = lib or _cabi.native()
        self._handle = None
        if output_type == "consensus":
            self._mask = OUTPUT_CONSENSUS
        elif output_type == "msa":
            self._mask = OUTPUT_MSA
        else:
            raise RuntimeError("Unknown output_type provided. Must be consensus/msa.")
        if stream is None:
            stream_ptr = None
        elif isinstance(stream, int):
            stream_ptr = ctypes.c_void_p(stream)
        elif hasattr(stream, "cuda_stream"):  # torch.cuda.Stream
            stream_ptr = ctypes.c_void_p(stream.cuda_stream)
        elif hasattr(stream, "stream"):  # genomeworks.cuda.CudaStream look-alike
            stream_ptr = ctypes.c_void_p(stream.stream)
        else:
            raise RuntimeError("Type for stream option must be CudaStream")

        if batch_size_c is not None:
            bs = batch_size_c
        else:
            # cudapoa.pyx:117-133: always the six-argument BatchSize constructor
            bs = BatchSizeC()
            self._lib.check(self._lib.batch_size_explicit(
                ctypes.byref(bs), max_sequence_size,
                2 * max_sequence_size if max_consensus_size is None else max_consensus_size,
                3 * max_sequence_size if max_nodes_per_window is None else max_nodes_per_window,
                4 * max_sequence_size if max_nodes_per_window_banded is None else max_nodes_per_window_banded,
                alignment_band_width, max_sequences_per_poa), "BatchSize")
        self.batch_size = bs
        handle = self._lib.create_batch(device_id, stream_ptr, int(max_gpu_mem), self._mask, ctypes.byref(bs),
                                        gap_score, mismatch_score, match_score, band_mode_value(cuda_banded_alignment))
        if not handle:
            msg = self._lib.error()
            if "negative" in msg or "should" in msg:
                raise ValueError(msg)
            raise RuntimeError(msg or "create_batch failed")
        self._handle = ctypes.c_void_p(handle)
        self._keepalive = None

    # -- lifetime -----------------------------------------------------------------------------------------
    def close(self):
        if self._handle is not None:
            self._lib.destroy_batch(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- pygenomeworks API --------------------------------------------------------------------------------
    def add_poa_group(self, poa):
        """Add one POA group (a list of str).  Returns (status, per-sequence status list)."""
        if not isinstance(poa, list):
            poa = [poa]
        if len(poa) < 1:
            raise RuntimeError("At least one sequence must be present in POA group")
        n = len(poa)
        byte_list = [seq.encode("utf-8") if isinstance(seq, str) else bytes(seq) for seq in poa]
        seqs = (ctypes.c_char_p * n)(*byte_list)
        lengths = (ctypes.c_int32 * n)(*[len(b) for b in byte_list])
        status = (ctypes.c_int32 * n)()
        n_status = ctypes.c_int32(0)
        rc = self._lib.add_poa_group(self._handle, seqs, None, lengths, n, status, ctypes.byref(n_status))
        self._lib.check(rc, "add_poa_group")
        return (rc, [status[i] for i in range(n_status.value)])

    @property
    def total_poas(self):
        return self._lib.check(self._lib.get_total_poas(self._handle), "get_total_poas")

    @property
    def batch_id(self):
        return self._lib.check(self._lib.batch_id(self._handle), "batch_id")

    def generate_poa(self):
        """Enqueue the partial order alignment of every group in the batch (asynchronous)."""
        self._lib.check(self._lib.generate_poa(self._handle), "generate_poa")

    def reset(self):
        self._lib.check(self._lib.reset(self._handle), "reset")

    def _statuses(self):
        n = self._lib.result_status_count(self._handle)
        return [self._lib.result_status(self._handle, i) for i in range(n)]

    def get_consensus(self):
        """Returns (list of consensus str, list of coverage lists, list of status)."""
        rc = self._lib.check(self._lib.get_consensus(self._handle), "get_consensus")
        if rc == output_type_unavailable:
            raise RuntimeError("Output type not requested during batch initialization")
        n = self._lib.result_count(self._handle)
        consensus, coverage = [], []
        seq = ctypes.c_void_p()
        cov = ctypes.c_void_p()
        for i in range(n):
            length = self._lib.result_consensus(self._handle, i, ctypes.byref(seq), ctypes.byref(cov))
            consensus.append(ctypes.string_at(seq.value, length).decode("utf-8") if length > 0 else "")
            if length > 0:
                coverage.append(list((ctypes.c_uint16 * length).from_address(cov.value)))
            else:
                coverage.append([])
        return (consensus, coverage, self._statuses())

    def get_msa(self):
        """Returns (list of MSAs, each a list of row str; list of status)."""
        rc = self._lib.check(self._lib.get_msa(self._handle), "get_msa")
        if rc == output_type_unavailable:
            raise RuntimeError("Output type not requested during batch initialization")
        n = self._lib.result_count(self._handle)
        out = []
        seq = ctypes.c_void_p()
        for i in range(n):
            rows = []
            for r in range(self._lib.result_msa_rows(self._handle, i)):
                length = self._lib.result_msa_row(self._handle, i, r, ctypes.byref(seq))
                rows.append(ctypes.string_at(seq.value, length).decode("utf-8"))
            out.append(rows)
        return (out, self._statuses())

    def get_graph_arrays(self):
        """get_graphs() as plain arrays: per group (labels bytes, [(src, dst, weight), ...]); and status list."""
        self._lib.check(self._lib.get_graphs(self._handle), "get_graphs")
        n = self._lib.result_count(self._handle)
        graphs = []
        for i in range(n):
            n_nodes = self._lib.result_graph_nodes(self._handle, i, None, 0)
            labels = ctypes.create_string_buffer(max(n_nodes, 1))
            self._lib.result_graph_nodes(self._handle, i, labels, n_nodes)
            n_edges = self._lib.result_graph_edges(self._handle, i, None, None, None, 0)
            src = np.zeros(max(n_edges, 1), dtype=np.int32)
            dst = np.zeros(max(n_edges, 1), dtype=np.int32)
            wgt = np.zeros(max(n_edges, 1), dtype=np.int32)
            self._lib.result_graph_edges(self._handle, i, src.ctypes.data, dst.ctypes.data, wgt.ctypes.data, n_edges)
            edges = [(int(src[k]), int(dst[k]), int(wgt[k])) for k in range(n_edges)]
            graphs.append((labels.raw[:n_nodes], edges))
        return (graphs, self._statuses())

    def get_graphs(self):
        """Returns (list of networkx.DiGraph with `label` node and `weight` edge attributes, status list)
        as cudapoa.pyx:258-292 does."""
        import networkx as nx

        graphs, status = self.get_graph_arrays()
        out = []
        for labels, edges in graphs:
            g = nx.DiGraph()
            for (s, d, w) in edges:
                g.add_edge(s, d, weight=w)
            attrs = {}
            for node in g.nodes:
                attrs[node] = {"label": chr(labels[node])}
            nx.set_node_attributes(g, attrs)
            out.append(g)
        return (out, status)

    # -- flat-buffer additions ------------------------------------------------------------------------------
    def add_poa_groups_flat(self, bases, read_lengths, window_n_reads, weights=None, want_seq_status=False):
        """Add many groups from flat numpy buffers (see cudapoa_b200_add_poa_groups_flat).
        Returns (status of the last add_poa_group, number of windows accepted[, per-read status array])."""
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        read_lengths = np.ascontiguousarray(read_lengths, dtype=np.int32)
        window_n_reads = np.ascontiguousarray(window_n_reads, dtype=np.int32)
        w_ptr = None
        if weights is not None:
            weights = np.ascontiguousarray(weights, dtype=np.int8)
            w_ptr = weights.ctypes.data
        seq_status = None
        s_ptr = None
        if want_seq_status:
            seq_status = np.full(read_lengths.shape[0], -1, dtype=np.int32)
            s_ptr = seq_status.ctypes.data
        n_added = ctypes.c_int32(0)
        rc = self._lib.add_poa_groups_flat(self._handle, bases.ctypes.data, w_ptr, read_lengths.ctypes.data,
                                           window_n_reads.ctypes.data, int(window_n_reads.shape[0]),
                                           ctypes.byref(n_added), s_ptr)
        self._lib.check(rc, "add_poa_groups_flat")
        if want_seq_status:
            return (rc, n_added.value, seq_status)
        return (rc, n_added.value)

    def fetch(self, kind="consensus"):
        """Run get_consensus/get_msa and leave the decoded outputs inside the handle (no Python objects)."""
        fn = self._lib.get_consensus if kind == "consensus" else self._lib.get_msa
        return self._lib.check(fn(self._handle), "get_" + kind)

    def result_digest(self):
        return int(self._lib.result_digest(self._handle))

    def result_total_chars(self):
        return int(self._lib.result_total_chars(self._handle))

    def result_statuses(self):
        return self._statuses()

    def stats(self):
        out = StatsC()
        if self._lib.get_stats(self._handle, ctypes.byref(out)) != 0:
            return None
        return {name: getattr(out, name) for name, _ in StatsC._fields_}

    @property
    def max_poas(self):
        return self._lib.max_poas(self._handle)
