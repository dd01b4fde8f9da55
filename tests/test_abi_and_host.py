"""CPU: the C-ABI library loads and exports every symbol include/cudapoa_b200.h declares; BatchSize semantics
(batch.hpp:75-127) through the boundary; window generator determinism.  No device work is started here."""
import ctypes
import os
import re

import numpy as np
import pytest

from claragenomicsanalysis_b200 import _cabi, build
from claragenomicsanalysis_b200.cudapoa import BatchSizeC, batch_size, status_to_str
from claragenomicsanalysis_b200.windows import WindowSet, generate_windows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _cabi.native()


def test_header_symbols_all_exported(lib):
    header = open(os.path.join(ROOT, "include", "cudapoa_b200.h")).read()
    declared = set(re.findall(r"CUDAPOA_CABI\((\w+)\)\s*\(", header))
    declared.discard("name")
    assert declared == set(_cabi.ABI_SYMBOLS)
    dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
    for name in sorted(declared):
        assert getattr(dll, "cudapoa_b200_" + name) is not None
    for extra in ("debug_rcp_table", "debug_dpx_peak", "debug_set_schedule", "bench_relaunch", "bench_relaunch_phases", "bench_relaunch_timed"):
        assert getattr(dll, "cudapoa_b200_" + extra) is not None
    assert lib.implementation() == b"b200-native"


def test_cpp_api_symbols_exported():
    """The C++ plugin interface itself (what a C++ caller links against): create_batch, Init, get_multi_batch_sizes."""
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", "-C", _cabi.NATIVE_LIB_PATH], stdout=subprocess.PIPE,
                         text=True).stdout
    for sym in ("claraparabricks::genomeworks::cudapoa::create_batch(",
                "claraparabricks::genomeworks::cudapoa::Init()",
                "claraparabricks::genomeworks::cudapoa::get_multi_batch_sizes("):
        assert sym in out, sym


def test_batch_size_default_ctor(lib):
    bs = batch_size(1024, 100, 256).as_dict()
    assert bs == dict(max_sequence_size=1024, max_consensus_size=2048, max_nodes_per_window=3072,
                      max_nodes_per_window_banded=4096, max_matrix_graph_dimension=3072,
                      max_matrix_graph_dimension_banded=4096, max_matrix_sequence_dimension=1024,
                      alignment_band_width=256, max_sequences_per_poa=100)
    # band width is rounded up to a multiple of 128 (batch.hpp:86-95)
    assert batch_size(500, 10, 200).alignment_band_width == 256
    assert batch_size(501, 10, 256).max_nodes_per_window == 1504  # align4(3 * 501)


def test_batch_size_validation(lib):
    out = BatchSizeC()
    assert lib.batch_size_default(ctypes.byref(out), -1, 10, 256) == -1
    assert "max_sequence_size cannot be negative" in lib.error()
    # six-argument constructor checks (batch.hpp:120-127)
    assert lib.batch_size_explicit(ctypes.byref(out), 1024, 2048, 512, 4096, 256, 10) == -1
    assert "max_nodes_per_window should be greater" in lib.error()
    assert lib.batch_size_explicit(ctypes.byref(out), 1024, 512, 3072, 4096, 256, 10) == -1
    assert "max_consensus_size should be greater" in lib.error()
    assert lib.batch_size_explicit(ctypes.byref(out), 128, 256, 384, 512, 256, 10) == -1
    assert "alignment_band_width should not be greater" in lib.error()
    assert lib.batch_size_explicit(ctypes.byref(out), 1024, 2048, 3072, 4096, 256, 10) == 0


def test_status_strings():
    assert status_to_str(0) == "success"
    assert status_to_str(1) == "exceeded_maximum_poas"
    assert status_to_str(9) == "generic_error"
    with pytest.raises(RuntimeError):
        status_to_str(42)


def test_create_batch_fails_loudly_without_a_device(lib):
    """No CPU fallback: without a GPU create_batch must not hand out a batch."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    # run in a child process: the reference behaviour on a CUDA failure is abort() (cudautils.hpp:77-88)
    import subprocess
    import sys

    code = ("import sys; sys.path.insert(0, %r);"
            "from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch;"
            "b = CudaPoaBatch(10, 256, 1 << 28); print('created')" % ROOT)
    res = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert res.returncode != 0
    assert "created" not in res.stdout


def test_window_generator_is_deterministic():
    a = generate_windows(3, start_index=7, template_length=120, n_reads=5)
    b = generate_windows(3, start_index=7, template_length=120, n_reads=5)
    assert np.array_equal(a.bases, b.bases) and np.array_equal(a.read_lengths, b.read_lengths)
    c = generate_windows(1, start_index=8, template_length=120, n_reads=5)
    assert a.window(1) == c.window(0)
    assert set(a.bases.tolist()) <= set(b"ACGT")
    w = generate_windows(2, template_length=80, n_reads=4, with_weights=True)
    assert w.weights.shape == w.bases.shape and w.weights.min() >= 0


def test_window_set_slice_tile_roundtrip():
    ws = generate_windows(5, template_length=60, n_reads=3)
    t = ws.tile(12)
    assert len(t) == 12 and t.window(7) == ws.window(2)
    s = t.slice(5, 9)
    assert [s.window(i) for i in range(4)] == [ws.window(i % 5) for i in range(5, 9)]
    r = WindowSet.from_lists([ws.window(i) for i in range(5)])
    assert np.array_equal(r.bases, ws.bases)
