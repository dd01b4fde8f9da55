"""Adaptive banding: an extension of the native engine with no counterpart in the reference
snapshot, hence no bit-exact reference oracle.  Its yardstick is the reference's *full* alignment: on windows whose
reads leave the diagonal by more than the static band follows, adaptive banding must agree with the full alignment at
least as often as the static band does; on ordinary windows it must not change the static band's result.  The CPU
oracle restates the widening rule independently (oracle/cudapoa_oracle.cpp) and the engine must match it bit for bit."""
import pytest

import cases
import oraclelib
import parity


@pytest.fixture(scope="module")
def native_lib():
    from claragenomicsanalysis_b200 import _cabi

    return _cabi.native()  # raises if the CUDA extension is missing: no fallback

KW = dict(max_seq=1024, max_reads=12)


def _oracle_adaptive(ws, band, **kw):
    p = oraclelib.make_params(band=band, adaptive=True, **kw)
    r = oraclelib.OracleRun(ws, p)
    return r, [r.consensus(w) for w in range(len(ws))]


def test_oracle_adaptive_recovers_full_alignment():
    ws = cases.structural_windows(12)
    static = parity.run_oracle(ws, band=128, **KW)
    full = parity.run_oracle(ws, banded=False, **KW)
    r, adaptive = _oracle_adaptive(ws, 128, **KW)
    agree_static = sum(a == b for a, b in zip(static["consensus"], full["consensus"]))
    agree_adaptive = sum(a[0] == b for a, b in zip(adaptive, full["consensus"]))
    assert r.widened > 0
    assert agree_static < len(ws), "the case must defeat the static band, or it tests nothing"
    assert agree_adaptive >= agree_static
    assert agree_adaptive >= len(ws) - 1


def test_oracle_adaptive_leaves_ordinary_windows_alone():
    ws = cases.case_windows("cfg2_banded16_consensus")
    static = parity.run_oracle(ws, 640, 30)
    r, adaptive = _oracle_adaptive(ws, 256, max_seq=640, max_reads=30)
    assert [a[0] for a in adaptive] == static["consensus"]
    assert [list(a[1]) for a in adaptive] == [list(c) for c in static["coverage"]]
    assert r.widened <= len(ws)  # a handful of low-scoring reads may be re-aligned; the result is the same


@pytest.fixture
def schedule(native_lib):
    """Process-wide schedule override (1: single launch, 2: one launch per phase and read) for one test."""
    def set_mode(mode):
        native_lib._dll.cudapoa_b200_debug_set_schedule(int(mode))
    yield set_mode
    native_lib._dll.cudapoa_b200_debug_set_schedule(0)


@pytest.mark.gpu
@pytest.mark.parametrize("mode,band", [(1, 128), (2, 128), (1, 256), (2, 256)])
def test_engine_adaptive_matches_oracle_and_full_alignment(native_lib, schedule, mode, band):
    """Both device schedules: the single launch re-aligns inside the window's loop, the phase-split schedule in
    poa_widen_kernel after the packed fill + traceback of the step (band 256) or the exact alignment (band 128)."""
    schedule(mode)
    ws = cases.structural_windows(12)
    got = parity.run_abi(native_lib, ws, band=band, banded="adaptive", **KW)
    r, want = _oracle_adaptive(ws, band, **KW)
    assert got["status"] == r.statuses()
    assert got["consensus"] == [w[0] for w in want]
    assert [list(c) for c in got["coverage"]] == [list(w[1]) for w in want]
    assert sum(s["widened"] for s in got["stats"]) == r.widened > 0
    if band != 128:
        return
    full = parity.run_abi(native_lib, ws, banded=False, **KW)
    static = parity.run_abi(native_lib, ws, band=128, banded=True, **KW)
    agree_static = sum(a == b for a, b in zip(static["consensus"], full["consensus"]))
    agree_adaptive = sum(a == b for a, b in zip(got["consensus"], full["consensus"]))
    assert agree_adaptive >= agree_static and agree_adaptive >= len(ws) - 1


@pytest.mark.gpu
def test_engine_adaptive_default_schedule(native_lib):
    ws = cases.structural_windows(12)
    got = parity.run_abi(native_lib, ws, band=128, banded="adaptive", **KW)
    r, want = _oracle_adaptive(ws, 128, **KW)
    assert got["status"] == r.statuses()
    assert got["consensus"] == [w[0] for w in want]
    assert [list(c) for c in got["coverage"]] == [list(w[1]) for w in want]
    assert sum(s["widened"] for s in got["stats"]) == r.widened > 0
    full = parity.run_abi(native_lib, ws, banded=False, **KW)
    static = parity.run_abi(native_lib, ws, band=128, banded=True, **KW)
    agree_static = sum(a == b for a, b in zip(static["consensus"], full["consensus"]))
    agree_adaptive = sum(a == b for a, b in zip(got["consensus"], full["consensus"]))
    assert agree_adaptive >= agree_static and agree_adaptive >= len(ws) - 1


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [1, 2])
def test_engine_adaptive_on_ordinary_windows(native_lib, schedule, mode):
    schedule(mode)
    ws = cases.case_windows("cfg2_banded16_consensus")
    got = parity.run_abi(native_lib, ws, 640, 30, banded="adaptive")
    static = parity.run_abi(native_lib, ws, 640, 30, banded=True)
    r, want = _oracle_adaptive(ws, 256, max_seq=640, max_reads=30)
    assert got["consensus"] == [w[0] for w in want] == static["consensus"]
    assert sum(s["widened"] for s in got["stats"]) == r.widened


@pytest.mark.gpu
def test_engine_adaptive_msa_and_int32(native_lib):
    ws = cases.case_windows("banded32_msa")
    e = cases.CASES["banded32_msa"]["engine"]
    got = parity.run_abi(native_lib, ws, e["max_seq"], e["max_reads"], band=e["band"], banded="adaptive", output="msa")
    p = oraclelib.make_params(max_seq=e["max_seq"], max_reads=e["max_reads"], band=e["band"], output="msa",
                              adaptive=True)
    r = oraclelib.OracleRun(ws, p)
    assert got["status"] == r.statuses()
    assert got["msa"] == [r.msa(w) for w in range(len(ws))]
