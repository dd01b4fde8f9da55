"""GPU: cudapoa_b200_consensus_noalloc (include/cudapoa_b200.h) -- the entry point that allocates nothing and keeps all
per-window state in a file-scope device array -- against the existing path (create_batch -> add_poa_group ->
generate_poa -> get_consensus) through the C ABI.  Integer work: outputs must be identical.

Collected last on purpose; passes alone in a fresh process and after the other tests.
Device buffers for inputs and outputs are the CALLER's business: the test makes them with torch, the entry point itself
never allocates."""
import ctypes

import numpy as np
import pytest

import parity
from claragenomicsanalysis_b200 import _cabi
from claragenomicsanalysis_b200.windows import generate_windows

SCRATCH_BYTES = 32 * 1024 * 1024  # CUDAPOA_B200_NOALLOC_SCRATCH_BYTES


_ENTRY = None


def _entry():
    """The entry point, bound once.  The library only exists after an nvcc build: skip (not error) where it is absent."""
    global _ENTRY
    if _ENTRY is None:
        import os

        if not os.path.exists(_cabi.NATIVE_LIB_PATH):
            pytest.skip("libcudapoa_b200.so has not been built on this machine")
        dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
        f = dll.cudapoa_b200_consensus_noalloc
        f.restype = ctypes.c_int32
        f.argtypes = ([ctypes.c_int32, ctypes.c_void_p] + [ctypes.c_void_p] * 3 + [ctypes.c_int32] * 4 +
                      [ctypes.c_int16] * 3 + [ctypes.c_void_p] * 4 + [ctypes.c_char_p, ctypes.c_int32])
        _ENTRY = f
    return _ENTRY


def test_noalloc_symbol_and_limits_without_gpu():
    """The size checks come before any CUDA call: refusals are testable on a CPU-only machine."""
    f = _entry()
    msg = ctypes.create_string_buffer(400)
    # 13 windows of BatchSize(640, 30, 256) do not fit 32 MiB
    rc = f(0, None, 1, 1, 1, 13, 30, 640, 256, -8, -6, 8, 1, 1, 1, 1, msg, 400)
    assert rc == -3
    text = msg.value.decode()
    assert "maximum is %d" % SCRATCH_BYTES in text and "13 windows" in text
    # a configuration the reference would run with 32-bit scores
    assert f(0, None, 1, 1, 1, 1, 10, 2048, 256, -8, -6, 8, 1, 1, 1, 1, msg, 400) == -2
    assert f(0, None, 1, 1, 1, 1, 10, -5, 256, -8, -6, 8, 1, 1, 1, 1, msg, 400) == -1
    assert f(0, None, None, 1, 1, 1, 10, 640, 256, -8, -6, 8, 1, 1, 1, 1, None, 0) == -1
    assert f(0, None, None, None, None, 0, 10, 640, 256, -8, -6, 8, None, None, None, None, None, 0) == 0


def _run_noalloc(f, torch, ws, max_reads, max_seq, band=256):
    n = len(ws)
    C = 2 * max_seq
    dev = torch.device("cuda:0")
    bases = torch.from_numpy(np.ascontiguousarray(ws.bases)).to(dev)
    lens = torch.from_numpy(np.ascontiguousarray(ws.read_lengths, dtype=np.int32)).to(dev)
    counts = torch.from_numpy(np.ascontiguousarray(ws.window_n_reads, dtype=np.int32)).to(dev)
    cons = torch.full((n, C), 0x7F, dtype=torch.uint8, device=dev)
    cov = torch.full((n, C), 0x7FFF, dtype=torch.int16, device=dev)
    length = torch.full((n,), -1, dtype=torch.int32, device=dev)
    status = torch.full((n,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    msg = ctypes.create_string_buffer(400)
    stream = torch.cuda.current_stream().cuda_stream
    rc = f(0, ctypes.c_void_p(stream), bases.data_ptr(), lens.data_ptr(), counts.data_ptr(), n, max_reads, max_seq, band,
           -8, -6, 8, cons.data_ptr(), cov.data_ptr(), length.data_ptr(), status.data_ptr(), msg, 400)
    assert rc == 0, msg.value.decode()
    torch.cuda.synchronize()
    cons_h = cons.cpu().numpy()
    cov_h = cov.cpu().numpy().view(np.uint16)
    len_h = length.cpu().numpy()
    out_c, out_v = [], []
    for w in range(n):
        k = int(len_h[w])
        out_c.append(cons_h[w, :k].tobytes().decode("ascii"))
        out_v.append([int(x) for x in cov_h[w, :k]])
        if k < C:
            assert cons_h[w, k] == 0  # NUL terminated
    return out_c, out_v, [int(s) for s in status.cpu().numpy()]


@pytest.mark.gpu
def test_noalloc_matches_batch_interface_on_two_calls_in_a_row():
    import torch

    lib = _cabi.native()  # raises if the CUDA extension is missing: no fallback
    f = _entry()
    # call 1: 12 windows of the benchmark's shape (the most BatchSize(640, 30, 256) allows); call 2: fewer, different,
    # smaller windows over the same scratch; call 3: the first set again (stale state of another shape underneath)
    first = generate_windows(12, start_index=31000, template_length=500, n_reads=30)
    second = generate_windows(7, start_index=32000, template_length=(120, 400), n_reads=9)
    for ws, max_reads, max_seq in ((first, 30, 640), (second, 9, 512), (first, 30, 640)):
        want = parity.run_abi(lib, ws, max_seq=max_seq, max_reads=max_reads, band=256, banded=True, output="consensus",
                              max_mem=2 << 30)
        got_c, got_v, got_s = _run_noalloc(f, torch, ws, max_reads, max_seq)
        assert got_s == [int(s) for s in want["status"]] == [0] * len(ws)
        assert got_c == want["consensus"]
        assert got_v == [list(map(int, c)) for c in want["coverage"]]
    # precondition violated on the device side: more reads than max_sequences_per_poa -> generic_error, length 0
    got_c, _, got_s = _run_noalloc(f, torch, second, 5, 512)
    assert got_s == [9] * len(second) and got_c == [""] * len(second)
