"""The engine's device code run on the CPU through the SIMT emulator in tests/emu (test infrastructure: g++ build of
csrc/engine.cu + csrc/c_api.cpp against stand-in CUDA headers, warps as fibers) and compared bit for bit with the
oracle and the reference goldens.  Catches kernel-logic regressions where no GPU is available; the GPU parity tests
(test_gpu_parity.py) remain the proof for the product library."""
import ctypes
import os
import sys

import pytest

import cases
import oraclelib
import parity
from claragenomicsanalysis_b200 import _cabi
from claragenomicsanalysis_b200.windows import generate_windows

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import build as emu_build  # noqa: E402


@pytest.fixture(scope="module")
def emu():
    os.environ["POA_EMU_RCP_TABLE"] = oraclelib.RCP_TABLE
    lib = _cabi.CabiLibrary(emu_build.build(), "cudapoa_b200_")
    lib.set_schedule = lib._dll.cudapoa_b200_debug_set_schedule
    lib.set_schedule.argtypes = [ctypes.c_int32]
    lib.set_schedule.restype = None
    yield lib
    lib.set_schedule(0)


def _check(lib, ws, eng, output="consensus", graphs=True):
    n = parity.run_abi(lib, ws, want_graphs=graphs, max_mem=1 << 30, **eng)
    o = parity.run_oracle(ws, want_graphs=graphs, **eng)
    assert parity.compare(n, o, output) == []
    if graphs:
        for w in range(len(ws)):
            assert n["graphs"][w] == o["graphs"][w], "graph %d" % w
    return n


@pytest.mark.parametrize("schedule", [1, 2])  # single launch, phase-split
def test_emulated_banded_consensus_matches_oracle(emu, schedule):
    emu.set_schedule(schedule)
    ws = generate_windows(6, start_index=42, template_length=200, n_reads=8)
    n = _check(emu, ws, dict(max_seq=256, max_reads=8, band=256, banded=True, output="consensus"))
    assert sum(s["exact_path"] for s in n["stats"]) == 0  # the packed path served every alignment


@pytest.mark.parametrize("schedule", [1, 2])
def test_emulated_msa_matches_oracle(emu, schedule):
    emu.set_schedule(schedule)
    ws = generate_windows(3, start_index=77, template_length=150, n_reads=6)
    _check(emu, ws, dict(max_seq=256, max_reads=6, band=256, banded=True, output="msa"), output="msa", graphs=False)


@pytest.mark.parametrize("trace", [4, 8])  # traceback by one warp / by one thread per window
def test_emulated_noisy_reads_phase_split(emu, trace):
    """25 % errors: bubbles, far predecessors (beyond the shared-memory ring), rows with several predecessors."""
    emu.set_schedule(2 + trace)
    ws = generate_windows(3, start_index=4000, template_length=300, n_reads=10, read_error=0.25, backbone_error=0.15)
    _check(emu, ws, dict(max_seq=448, max_reads=10, band=256, banded=True, output="consensus"))


@pytest.mark.parametrize("length,reads", [(900, 4), (1600, 4), (2600, 4)])
def test_emulated_long_window_rebased_rows(emu, length, reads):
    """Reference int32-score range (L >= 1490): the packed fill stores rows relative to a per-row offset and keeps a
    sliding 512-column window of the read's profile (2 600 bases: the window is refilled eight times per read).  Reads
    of more than 768 bases take this fill too, whatever their score range (900)."""
    emu.set_schedule(2)
    ws = generate_windows(1, start_index=7000, template_length=length, n_reads=reads)
    n = _check(emu, ws, dict(max_seq=3200, max_reads=4, band=256, banded=True, output="consensus"))
    assert sum(s["exact_path"] for s in n["stats"]) == 0


@pytest.mark.parametrize("gen,eng", [
    # reads shorter than the band: the match-bit profile's padding, bands that never move
    (dict(n=3, start_index=9500, template_length=90, n_reads=6, read_error=0.15),
     dict(max_seq=256, max_reads=6, band=256, banded=True, output="consensus")),
    # full alignment of 700-base reads: three column chunks per row, traceback across chunk edges
    (dict(n=2, start_index=9600, template_length=700, n_reads=5, read_error=0.12),
     dict(max_seq=768, max_reads=5, band=256, banded=False, output="consensus")),
    # windows of 60 .. 620 bases in one batch: profile strides and band placement of every size class
    (dict(n=6, start_index=9700, template_length=(60, 620), n_reads=8, read_error=0.15),
     dict(max_seq=704, max_reads=8, band=256, banded=True, output="consensus")),
])
def test_emulated_packed_fill_edge_shapes(emu, gen, eng):
    """Shapes at the edges of the packed fill's data layout (tools/emu_check.py runs the same and more at larger counts)."""
    gen = dict(gen)
    ws  = generate_windows(gen.pop("n"), **gen)
    for mode in (2 + 4, 2 + 8):  # traceback by one warp / one thread per window
        emu.set_schedule(mode)
        n = _check(emu, ws, eng, graphs=False)
        assert sum(s["exact_path"] for s in n["stats"]) == 0


def test_emulated_cfg2_window_matches_reference_golden(emu):
    """One full-size cfg 2 window (500 bp x 30 reads) against the golden produced by the unmodified reference."""
    emu.set_schedule(2)
    g = parity.load_golden("reference_cfg2_banded16_consensus.json")
    if g is None:
        pytest.skip("golden file missing")
    c = cases.CASES["cfg2_banded16_consensus"]
    ws = cases.case_windows("cfg2_banded16_consensus").slice(0, 1)
    n = parity.run_abi(emu, ws, max_mem=1 << 30, **c["engine"])
    assert parity.digests(n, "consensus") == g["digest"][:1]


def test_emulated_pipeline_worker_pool(emu):
    """The window pipeline (csrc/pipeline.cpp: persistent workers, cost-ordered guided chunks) over the emulated engine:
    mixed window sizes, three workers, two runs on the same pool; results indexed by window and equal to the oracle."""
    from claragenomicsanalysis_b200.pipeline import WindowPipeline

    emu.set_schedule(0)
    ws = generate_windows(14, start_index=900, template_length=(60, 240), n_reads_range=(3, 7))
    eng = dict(max_seq=320, max_reads=8, band=256, banded=True, output="consensus")
    o = parity.run_oracle(ws, **eng)
    p = WindowPipeline(8, 320, 1 << 28, devices=(0,), workers_per_device=3, output_type="consensus", lib=emu)
    try:
        assert p.workers == 3
        for chunk in (4, 0):
            p.run(ws, chunk_windows=chunk)
            assert p.statuses() == [int(s) for s in o["status"]]
            for w in range(len(ws)):
                assert p.consensus(w) == (o["consensus"][w], list(o["coverage"][w])), "window %d" % w
            assert p.stats()["windows"] == len(ws)
    finally:
        p.close()



@pytest.mark.parametrize("length,reads", [(300, 6), (520, 5)])
def test_emulated_full_alignment_packed_chunks(emu, length, reads):
    """Full alignment on the phase-split schedule: the packed fill runs the matrix in column chunks of 256 cells
    (two and three chunks here), the traceback follows the codes across chunk edges; nothing goes to the exact path."""
    emu.set_schedule(2)
    ws = generate_windows(2, start_index=8800, template_length=length, n_reads=reads)
    n = _check(emu, ws, dict(max_seq=640, max_reads=8, band=256, banded=False, output="consensus"))
    assert sum(s["exact_path"] for s in n["stats"]) == 0
