"""ctypes binding of the cudapoa C ABI declared in include/cudapoa_b200.h.

The product loads exactly one library, ``claragenomicsanalysis_b200/lib/libcudapoa_b200.so`` (built in-tree by
``build.py``), and fails loudly when it is missing: there is no CPU or Python fallback for the POA path.

``CabiLibrary`` takes the library path and symbol prefix as arguments because the same ABI is also compiled over the
reference implementation for the parity tests (prefix ``cudapoa_ref_``); nothing in this package points at it.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
NATIVE_LIB_PATH = os.path.join(_HERE, "lib", "libcudapoa_b200.so")

STATUS_NAMES = (
    "success",
    "exceeded_maximum_poas",
    "exceeded_maximum_sequence_size",
    "exceeded_maximum_sequences_per_poa",
    "node_count_exceeded_maximum_graph_size",
    "edge_count_exceeded_maximum_graph_size",
    "seq_len_exceeded_maximum_nodes_per_window",
    "loop_count_exceeded_upper_bound",
    "output_type_unavailable",
    "generic_error",
)

OUTPUT_CONSENSUS = 1
OUTPUT_MSA = 2


class BatchSizeC(ctypes.Structure):
    """cudapoa_b200_batch_size == cudapoa::BatchSize field order (batch.hpp:53-73)."""

    _fields_ = [
        ("max_sequence_size", ctypes.c_int32),
        ("max_consensus_size", ctypes.c_int32),
        ("max_nodes_per_window", ctypes.c_int32),
        ("max_nodes_per_window_banded", ctypes.c_int32),
        ("max_matrix_graph_dimension", ctypes.c_int32),
        ("max_matrix_graph_dimension_banded", ctypes.c_int32),
        ("max_matrix_sequence_dimension", ctypes.c_int32),
        ("alignment_band_width", ctypes.c_int32),
        ("max_sequences_per_poa", ctypes.c_int32),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


class StatsC(ctypes.Structure):
    _fields_ = [
        ("windows", ctypes.c_int64),
        ("reads_aligned", ctypes.c_int64),
        ("dp_cells", ctypes.c_int64),
        ("graph_rows", ctypes.c_int64),
        ("exact_path", ctypes.c_int64),
        ("h2d_bytes", ctypes.c_int64),
        ("d2h_bytes", ctypes.c_int64),
        ("kernel_ms", ctypes.c_double),
        ("graph_kernel_ms", ctypes.c_double),
        ("widened", ctypes.c_int64),
        ("kernel_launches", ctypes.c_int64),
    ]


# name -> (restype, argtypes); every symbol include/cudapoa_b200.h declares.
_P = ctypes.POINTER
_SIGNATURES = {
    "last_error": (ctypes.c_char_p, []),
    "implementation": (ctypes.c_char_p, []),
    "init": (ctypes.c_int32, []),
    "batch_size_default": (ctypes.c_int32, [_P(BatchSizeC), ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "batch_size_explicit": (ctypes.c_int32, [_P(BatchSizeC)] + [ctypes.c_int32] * 6),
    "create_batch": (ctypes.c_void_p, [ctypes.c_int32, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int8,
                                       _P(BatchSizeC), ctypes.c_int16, ctypes.c_int16, ctypes.c_int16,
                                       ctypes.c_int32]),
    "destroy_batch": (None, [ctypes.c_void_p]),
    "add_poa_group": (ctypes.c_int32, [ctypes.c_void_p, _P(ctypes.c_char_p), _P(ctypes.c_void_p),
                                       _P(ctypes.c_int32), ctypes.c_int32, _P(ctypes.c_int32), _P(ctypes.c_int32)]),
    "add_poa_groups_flat": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_int32, _P(ctypes.c_int32), ctypes.c_void_p]),
    "get_total_poas": (ctypes.c_int32, [ctypes.c_void_p]),
    "generate_poa": (ctypes.c_int32, [ctypes.c_void_p]),
    "batch_id": (ctypes.c_int32, [ctypes.c_void_p]),
    "reset": (ctypes.c_int32, [ctypes.c_void_p]),
    "get_consensus": (ctypes.c_int32, [ctypes.c_void_p]),
    "get_msa": (ctypes.c_int32, [ctypes.c_void_p]),
    "get_graphs": (ctypes.c_int32, [ctypes.c_void_p]),
    "result_count": (ctypes.c_int32, [ctypes.c_void_p]),
    "result_status_count": (ctypes.c_int32, [ctypes.c_void_p]),
    "result_status": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32]),
    "result_consensus": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, _P(ctypes.c_void_p),
                                          _P(ctypes.c_void_p)]),
    "result_msa_rows": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32]),
    "result_msa_row": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, _P(ctypes.c_void_p)]),
    "result_graph_nodes": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32]),
    "result_graph_edges": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_int32]),
    "result_total_chars": (ctypes.c_int64, [ctypes.c_void_p]),
    "result_digest": (ctypes.c_uint64, [ctypes.c_void_p]),
    "get_multi_batch_sizes": (ctypes.c_int32, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                               ctypes.c_int32, ctypes.c_int32, ctypes.c_float, ctypes.c_int32,
                                               ctypes.c_int32, ctypes.c_int32, _P(BatchSizeC), ctypes.c_int32,
                                               ctypes.c_void_p]),
    "get_stats": (ctypes.c_int32, [ctypes.c_void_p, _P(StatsC)]),
    "max_poas": (ctypes.c_int32, [ctypes.c_void_p]),
}

ABI_SYMBOLS = tuple(sorted(_SIGNATURES))


class CabiError(RuntimeError):
    pass


class CabiLibrary:
    """One loaded library exposing the cudapoa C ABI under ``prefix``."""

    def __init__(self, path, prefix):
        if not os.path.exists(path):
            raise CabiError(
                "cudapoa C-ABI library not found: %s. Build it with `python -m claragenomicsanalysis_b200.build` "
                "(nvcc, sm_100a). There is no CPU fallback for this path." % path)
        self.path = path
        self.prefix = prefix
        self._dll = ctypes.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(self._dll, prefix + name)  # AttributeError if the symbol is missing: loud by design
            fn.restype = restype
            fn.argtypes = argtypes
            setattr(self, name, fn)

    def error(self):
        msg = self.last_error()
        return msg.decode("utf-8", "replace") if msg else ""

    def check(self, rc, what):
        """Translate the boundary's exception codes back into the Python exceptions Cython's `except +` raises."""
        if rc == -1:
            raise ValueError("%s: %s" % (what, self.error()))  # std::invalid_argument
        if rc in (-2, -3):
            raise RuntimeError("%s: %s" % (what, self.error()))  # std::runtime_error / other
        return rc


_native = None


def native():
    """The product library (libcudapoa_b200.so).  Raises if it has not been built."""
    global _native
    if _native is None:
        _native = CabiLibrary(NATIVE_LIB_PATH, "cudapoa_b200_")
    return _native
