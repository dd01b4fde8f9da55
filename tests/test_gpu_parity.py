"""GPU: the native engine, through the C ABI, against (1) the CPU oracle on the same seeded windows, (2) the golden
outputs of the reference cudapoa, (3) the live reference build when oracle/_ref/libcudapoa_ref.so travelled, and
(4) size-independent properties at larger sizes.  Bit-exact everywhere: all arithmetic is integer."""
import random

import numpy as np
import pytest

import oraclelib
import parity
from cases import CASES, case_windows
from claragenomicsanalysis_b200 import _cabi
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch
from claragenomicsanalysis_b200.windows import WindowSet, generate_windows

pytestmark = pytest.mark.gpu

MEM = 4 << 30


@pytest.fixture(scope="module")
def lib():
    return _cabi.native()  # raises if the CUDA extension is missing: no fallback


@pytest.mark.parametrize("name", sorted(CASES))
def test_native_matches_oracle_and_reference_golden(lib, name):
    case = CASES[name]
    eng = case["engine"]
    output = eng["output"]
    ws = case_windows(name)
    n = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    assert parity.compare(n, o, output) == []
    for w in range(len(ws)):
        if n["status"][w] == 0:
            assert n["graphs"][w][0] == o["graphs"][w][0]
            assert n["graphs"][w][1] == o["graphs"][w][1]
    g = parity.load_golden("reference_%s.json" % name)
    assert g is not None
    assert [int(s) for s in n["status"]] == g["status"]
    assert parity.digests(n, output) == g["digest"]
    assert n["seq_status"] == g["seq_status"]
    stats = n["stats"][0]
    assert stats["dp_cells"] == o["dp_cells"]


@pytest.mark.parametrize("name", ["cfg2_banded16_consensus", "cfg3_banded16_msa", "full16_consensus",
                                  "banded32_consensus", "cfg4_banded32_real", "cfg5b_mixed", "banded_size32",
                                  "banded_size32_msa"])
def test_native_matches_live_reference(lib, name):
    if not oraclelib.have_ref_cudapoa():
        pytest.skip("reference build not present on this box")
    case = CASES[name]
    ws = case_windows(name)
    n = parity.run_abi(lib, ws, max_mem=MEM, **case["engine"])
    r = parity.run_abi(parity.ref_lib(), ws, max_mem=MEM, **case["engine"])
    assert parity.compare(n, r, case["engine"]["output"]) == []


# ---- the two device schedules of a banded batch: one launch per window set, or one launch per phase and read ----------
# (batches with int32 node ids always take the single launch: nothing to compare)
BANDED_CASES = sorted(n for n in CASES if CASES[n]["engine"].get("banded", True) and CASES[n]["engine"]["max_seq"] < 8192)


@pytest.fixture
def schedule(lib):
    """Sets the process-wide schedule override for one test and restores the default afterwards."""
    def set_mode(mode):
        lib._dll.cudapoa_b200_debug_set_schedule(int(mode))
    yield set_mode
    set_mode(0)


@pytest.mark.parametrize("name", BANDED_CASES)
def test_single_launch_and_phase_split_agree(lib, schedule, name):
    """poa_window_kernel (mode 1) and poa_fill_kernel + poa_graph_kernel (mode 2) run the same device functions in the
    same order per window: statuses, outputs, graphs, per-sequence statuses and work counters must be identical, and
    both must equal the oracle.  The default mode is covered by every other test of this file."""
    case = CASES[name]
    eng = case["engine"]
    ws = case_windows(name)
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    got = {}
    for mode in (1, 2):
        schedule(mode)
        got[mode] = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
        assert parity.compare(got[mode], o, eng["output"]) == []
        launches = sum(s["kernel_launches"] for s in got[mode]["stats"])
        assert (launches == len(got[mode]["stats"])) == (mode == 1)
    assert got[1]["status"] == got[2]["status"]
    assert got[1]["graphs"] == got[2]["graphs"]
    assert got[1]["seq_status"] == got[2]["seq_status"]
    for key in ("dp_cells", "graph_rows", "exact_path", "reads_aligned"):
        assert [s[key] for s in got[1]["stats"]] == [s[key] for s in got[2]["stats"]]
    # the phase-split schedule's two traceback kernels (one warp / one thread per window; the second serves large batches)
    for trace in (4, 8):
        schedule(2 + trace)
        t = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
        assert parity.compare(t, o, eng["output"]) == []
        assert t["graphs"] == got[2]["graphs"]


FULL_SPLIT_CASES = sorted(n for n in CASES if not CASES[n]["engine"].get("banded", True) and CASES[n]["engine"]["max_seq"] <= 768)


@pytest.mark.parametrize("name", FULL_SPLIT_CASES)
def test_full_alignment_packed_path_matches_exact_path_and_oracle(lib, schedule, name):
    """Full alignment (the reference's default mode): the single launch aligns with the exact-arithmetic path, the
    phase-split schedule with the packed fill run in column chunks of 256 cells (fill2_impl<.., FULL>) and the code
    traceback; both must equal the oracle bit for bit, and the packed path must decline nothing."""
    case = CASES[name]
    eng = case["engine"]
    ws = case_windows(name)
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    got = {}
    for mode in (1, 2):
        schedule(mode)
        got[mode] = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
        assert parity.compare(got[mode], o, eng["output"]) == []
    assert got[1]["graphs"] == got[2]["graphs"] == o["graphs"]
    assert got[1]["seq_status"] == got[2]["seq_status"]
    assert sum(s["exact_path"] for s in got[2]["stats"]) == 0
    assert sum(s["exact_path"] for s in got[1]["stats"]) == sum(s["reads_aligned"] for s in got[1]["stats"])
    assert [s["dp_cells"] for s in got[1]["stats"]] == [s["dp_cells"] for s in got[2]["stats"]]


def test_full_alignment_cfg2_windows_phase_split(lib, schedule):
    """cfg 2 windows (500 bp x 30 reads: three column chunks per row) in full-alignment mode through the phase-split
    schedule against the oracle, and -- when the reference build is present -- against the live reference."""
    ws = generate_windows(24, start_index=31000, template_length=500, n_reads=30)
    eng = dict(max_seq=640, max_reads=30, band=256, banded=False, output="consensus")
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    for mode in (2 + 4, 2 + 8):  # traceback across chunk edges by one warp / by one thread per window
        schedule(mode)
        got = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
        assert parity.compare(got, o, "consensus") == []
        assert got["graphs"] == o["graphs"]
        assert sum(s["exact_path"] for s in got["stats"]) == 0


@pytest.mark.parametrize("output", ["consensus", "msa"])
def test_ragged_batch_both_schedules(lib, schedule, output):
    """Windows of one batch finish at different launches of the phase-split schedule: read counts 1 .. 14 mixed with
    lengths 60 .. 600, a window whose graph outgrows max_nodes (status 4) and one that only has a backbone.  Both
    schedules must reproduce the oracle window by window."""
    rng = random.Random(20261105)
    pool = generate_windows(40, start_index=15000, template_length=(60, 600), n_reads=14)
    wins = []
    for i in range(len(pool)):
        reads = pool.window(i)
        wins.append(reads[:rng.choice([1, 2, 3, 5, 9, 14])])
    # reads over disjoint alphabets: every base becomes a new node, the node limit is reached before the last read
    junk = ["".join(rng.choice("ACGTEFHIJKLMNOPQRSUVWXYZ"[4 * (k % 6):4 * (k % 6) + 4]) for _ in range(500)) for k in range(12)]
    wins.insert(7, junk)
    ws = WindowSet.from_lists(wins)
    eng = dict(max_seq=640, max_reads=14, band=256, banded=True, output=output)
    o = parity.run_oracle(ws, want_graphs=True, **eng)
    assert len(set(o["status"])) > 1  # the case does contain a failing window
    for mode in (1, 2):
        schedule(mode)
        n = parity.run_abi(lib, ws, want_graphs=True, max_mem=MEM, **eng)
        assert parity.compare(n, o, output) == []
        assert [int(s) for s in n["status"]] == [int(s) for s in o["status"]]


def test_phase_split_many_windows_property(lib, schedule):
    """Size-independent property at a size the oracle does not reach: 4096 windows, every 512th input repeated, give
    identical outputs for identical inputs whatever warp, launch class or schedule handled them."""
    base = generate_windows(512, start_index=7000, template_length=400, n_reads=24)
    ws = base.tile(4096)
    eng = dict(max_seq=512, max_reads=24, band=256, banded=True, output="consensus")
    res = {}
    for mode in (1, 2, 2 + 8):  # 2 + 8: the phase-split schedule with the thread-per-window traceback
        schedule(mode)
        res[mode] = parity.run_abi(lib, ws, max_mem=16 << 30, **eng)
        r = res[mode]
        assert r["status"] == [0] * 4096
        for w in range(512, 4096):
            assert r["consensus"][w] == r["consensus"][w % 512]
            assert list(r["coverage"][w]) == list(r["coverage"][w % 512])
    assert res[1]["consensus"] == res[2]["consensus"] == res[10]["consensus"]
    assert [list(c) for c in res[1]["coverage"]] == [list(c) for c in res[2]["coverage"]] == [list(c) for c in res[10]["coverage"]]


# ---- Batch API behaviour, after cudapoa/tests/Test_CudapoaBatch.cu:66-199 -----------------------------------------
def test_initialize_and_zero_memory(lib):
    b = CudaPoaBatch(5, 1024, MEM, lib=lib)
    assert b.total_poas == 0
    b.close()
    with pytest.raises(RuntimeError):
        CudaPoaBatch(5, 1024, 0, lib=lib)


def test_add_and_reset_counts(lib):
    b = CudaPoaBatch(5, 1024, MEM, lib=lib)
    ws = WindowSet.from_lists([[""]])  # the all-zero Entry{} of Test_CudapoaBatch.cu:80-93
    rc, added = b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
    assert rc == 0 and added == 1 and b.total_poas == 1
    b.reset()
    assert b.total_poas == 0
    b.close()


def test_per_sequence_status_codes(lib):
    b = CudaPoaBatch(10, 1024, MEM, lib=lib)
    rc, st = b.add_poa_group(["A" * 20] * 11)
    assert rc == 0 and b.total_poas == 1
    assert st[:10] == [0] * 10 and st[10] == 3  # exceeded_maximum_sequences_per_poa
    rc, st = b.add_poa_group(["A" * 1024, "A" * 1025])
    assert rc == 0 and st == [0, 2]  # exceeded_maximum_sequence_size
    b.close()


def test_identical_sequences_give_themselves(lib):
    b = CudaPoaBatch(10, 1024, MEM, lib=lib)
    seq = "A" * 1023
    ws = WindowSet.from_lists([[seq, seq, seq]], weights=[[np.ones(1023, np.int8)] * 3])
    b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads, weights=ws.weights)
    b.generate_poa()
    cons, cov, st = b.get_consensus()
    assert st == [0] and cons == [seq] and cov[0] == [3] * 1023
    b.close()


def test_negative_weight_is_invalid_argument(lib):
    b = CudaPoaBatch(10, 1024, MEM, lib=lib)
    ws = WindowSet.from_lists([["ACGT", "ACGT"]], weights=[[np.array([1, 1, 1, 1], np.int8),
                                                              np.array([1, -1, 1, 1], np.int8)]])
    with pytest.raises(ValueError):
        b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads, weights=ws.weights)
    b.close()


def test_output_type_unavailable(lib):
    b = CudaPoaBatch(10, 256, MEM, output_type="msa", lib=lib)
    b.add_poa_group(["ACGTACGT", "ACGAACGT"])
    b.generate_poa()
    with pytest.raises(RuntimeError):
        b.get_consensus()
    msa, st = b.get_msa()
    assert st == [0] and len(msa[0]) == 2
    b.close()
    with pytest.raises(RuntimeError):
        CudaPoaBatch(10, 256, MEM, output_type="error_input", lib=lib)


def test_batch_full_returns_exceeded_maximum_poas(lib):
    bs_mem = 64 << 20
    b = CudaPoaBatch(4, 256, bs_mem, cuda_banded_alignment=True, lib=lib)
    cap = b.max_poas
    assert cap >= 1
    ws = generate_windows(cap + 3, template_length=100, n_reads=4)
    rc, added = b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads)
    assert rc == 1 and added == cap and b.total_poas == cap  # exceeded_maximum_poas
    b.close()


# ---- pygenomeworks binding behaviour, after pygenomeworks/test/test_cudapoa_bindings.py ----------------------------
@pytest.mark.parametrize("banded", [False, True])
def test_simple_batches(lib, banded):
    b = CudaPoaBatch(10, 1024, MEM, cuda_banded_alignment=banded, lib=lib)
    b.add_poa_group(["ACTGACTG", "ACTTACTG", "ACGGACTG", "ATCGACTG"])
    b.add_poa_group(["ACTGAC", "ACTTAC", "ACGGAC", "ATCGAC"])
    b.generate_poa()
    cons, cov, st = b.get_consensus()
    assert len(cons) == 2 and b.total_poas == 2 and st == [0, 0]
    b.reset()
    assert b.total_poas == 0
    b.close()


def test_graph_of_three_reads(lib):
    b = CudaPoaBatch(10, 1024, MEM, lib=lib)
    b.add_poa_group(["ACTGACTG", "ACTTACTG", "ACTCACTG"])
    b.generate_poa()
    b.get_consensus()
    graphs, st = b.get_graphs()
    assert len(graphs) == 1
    assert graphs[0].number_of_nodes() == 10 and graphs[0].number_of_edges() == 11
    b.close()


def test_complex_batch_recovers_truth(lib):
    random.seed(2)
    ref = "".join(random.choice("ACGT") for _ in range(500))
    reads = ["".join(r if random.random() > 0.02 else random.choice("ACGT") for r in ref) for _ in range(100)]
    b = CudaPoaBatch(1000, 1024, MEM, lib=lib)
    b.add_poa_group(reads)
    b.generate_poa()
    cons, cov, st = b.get_consensus()
    assert st == [0] and cons[0] == ref
    b.close()


# ---- MSA, after cudapoa/tests/Test_CudapoaGenerateMSA2.cu ---------------------------------------------------------
def test_msa_rows_spell_their_reads(lib):
    ws = generate_windows(1, start_index=31, template_length=50, n_reads=500, read_error=0.2)
    n = parity.run_abi(lib, ws, max_seq=1024, max_reads=500, banded=False, output="msa", max_mem=8 << 30)
    assert n["status"] == [0]
    rows = n["msa"][0]
    assert len(rows) == 500 and len(set(len(r) for r in rows)) == 1
    assert [r.replace("-", "") for r in rows] == ws.window(0)


def test_msa_overflow_status(lib):
    # max_consensus_size == max_sequence_size, backbone one short of it (Test_CudapoaGenerateMSA2.cu:127-158)
    bs = CudaPoaBatch(10, 1024, MEM, output_type="msa", lib=lib).batch_size
    bs.max_consensus_size = bs.max_sequence_size
    ws = generate_windows(1, start_index=77, template_length=1010, n_reads=10, backbone_error=0.0, read_error=0.15)
    n = parity.run_abi(lib, ws, max_seq=1024, max_reads=10, banded=False, output="msa", max_mem=MEM, batch_size_c=bs)
    o = parity.run_oracle(ws, max_seq=1024, max_reads=10, banded=False, output="msa", batch_size=bs)
    assert n["status"] == o["status"] == [2]  # exceeded_maximum_sequence_size


def test_spoa_accurate_mode_matches_spoa_msa(lib, monkeypatch):
    if not oraclelib.have_ref_spoa():
        pytest.skip("reference SPOA not built")
    monkeypatch.setenv("CUDAPOA_B200_SPOA_ACCURATE", "1")
    ws = generate_windows(4, start_index=500, template_length=120, n_reads=12)
    n = parity.run_abi(lib, ws, max_seq=256, max_reads=12, banded=False, output="msa", max_mem=MEM)
    o = parity.run_oracle(ws, max_seq=256, max_reads=12, banded=False, output="msa", spoa_accurate=True)
    assert parity.compare(n, o, "msa") == []
    sp = oraclelib.SpoaRun(ws, msa=True)
    assert [n["msa"][w] for w in range(4)] == [sp.msa(w) for w in range(4)]


# ---- properties at larger sizes ------------------------------------------------------------------------------------
def test_large_batch_properties(lib):
    """2048 cfg-2 windows: all succeed, results do not depend on how the windows are split into batches or on
    where a window sits in the batch, and a second run is identical."""
    base = generate_windows(256, template_length=500, n_reads=30)
    ws = base.tile(2048)
    eng = dict(max_seq=640, max_reads=30, band=256, banded=True, output="consensus")
    a = parity.run_abi(lib, ws, max_mem=24 << 30, **eng)
    assert set(a["status"]) == {0}
    da = parity.digests(a, "consensus")
    assert da[:256] * 8 == da  # tiled windows give tiled results
    b = parity.run_abi(lib, ws.slice(0, 600), max_mem=1 << 30, **eng)  # forces several batch rounds
    assert parity.digests(b, "consensus") == da[:600]
    o = parity.run_oracle(base.slice(0, 8), **eng)
    assert parity.digests(o, "consensus") == da[:8]
    lens = [len(c) for c in a["consensus"][:256]]
    assert 450 <= min(lens) and max(lens) <= 560
    assert all(max(cov) <= 30 for cov in a["coverage"][:256])


# ---- window pipeline (stream scheduler / sharding layer) ------------------------------------------------------------
@pytest.mark.parametrize("output", ["consensus", "msa"])
def test_pipeline_matches_single_batch_and_oracle(lib, output):
    from claragenomicsanalysis_b200.pipeline import WindowPipeline

    ws = generate_windows(96, start_index=12000, template_length=(150, 400), n_reads=9)
    eng = dict(max_seq=512, max_reads=9, band=256, banded=True, output=output)
    single = parity.run_abi(lib, ws, max_mem=MEM, **eng)
    o = parity.run_oracle(ws.slice(0, 12), **eng)
    # two "devices" (the same GPU twice) x two workers, chunks of 7 windows: many rounds, every worker busy
    p = WindowPipeline(9, 512, 1 << 30, devices=(0, 0), workers_per_device=2, output_type=output)
    assert p.workers == 4
    p.run(ws, chunk_windows=7)
    assert p.statuses() == [int(s) for s in single["status"]]
    for w in range(len(ws)):
        if output == "consensus":
            assert p.consensus(w) == (single["consensus"][w], list(single["coverage"][w]))
        else:
            assert p.msa(w) == list(single["msa"][w])
    for w in range(12):
        if output == "consensus":
            assert p.consensus(w) == (o["consensus"][w], list(o["coverage"][w]))
        else:
            assert p.msa(w) == o["msa"][w]
    st = p.stats()
    assert st["windows"] == 96 and st["reads_aligned"] == 96 * 8
    d1 = p.digest()
    p.run(ws, chunk_windows=0)  # automatic chunking, second run on the same pipeline
    assert p.digest() == d1
    p.close()


# ---- get_multi_batch_sizes (cudapoa/include/.../utils.hpp:48-58, cudapoa/src/utils.cu:24-138) ------------------------
def _multi_batch(lib, read_lengths, window_n_reads, banded=True, msa=False, band=256, quota=0.9):
    import ctypes

    n = len(window_n_reads)
    out = (_cabi.BatchSizeC * 64)()
    w2b = np.full(n, -2, dtype=np.int32)
    rl = np.ascontiguousarray(read_lengths, dtype=np.int32)
    wn = np.ascontiguousarray(window_n_reads, dtype=np.int32)
    k = lib.get_multi_batch_sizes(rl.ctypes.data, wn.ctypes.data, n, int(banded), int(msa), band, ctypes.c_float(quota),
                                  -6, -8, 8, out, 64, w2b.ctypes.data)
    assert k >= 0, lib.error()
    return [out[i].as_dict() for i in range(k)], w2b


def test_get_multi_batch_sizes_partitions_and_fits(lib):
    """Every group lands in exactly one batch, fits that batch's BatchSize (longest read and read count), batches come
    out largest windows first, and a batch created from its BatchSize accepts its groups."""
    ws = generate_windows(400, start_index=50000, length_choices=[100, 250, 500, 1000, 2000],
                          length_probs=[0.2, 0.2, 0.3, 0.2, 0.1], n_reads_range=(2, 40))
    sizes, w2b = _multi_batch(lib, ws.read_lengths, ws.window_n_reads)
    assert len(sizes) >= 1 and w2b.min() >= 0 and w2b.max() == len(sizes) - 1
    assert set(w2b.tolist()) == set(range(len(sizes)))
    off = np.concatenate([[0], np.cumsum(ws.window_n_reads)])
    for w in range(len(ws)):
        bs = sizes[w2b[w]]
        assert int(ws.read_lengths[off[w]:off[w + 1]].max()) <= bs["max_sequence_size"]
        assert int(ws.window_n_reads[w]) <= bs["max_sequences_per_poa"]
    longest = [s["max_sequence_size"] for s in sizes]
    assert longest == sorted(longest, reverse=True)
    # the first batch really takes its groups
    members = [w for w in range(len(ws)) if w2b[w] == 0][:8]
    sub = WindowSet.concat([ws.slice(w, w + 1) for w in members])
    r = parity.run_abi(lib, sub, max_seq=sizes[0]["max_sequence_size"], max_reads=sizes[0]["max_sequences_per_poa"],
                       max_mem=8 << 30)
    assert r["status"] == [0] * len(members) and set(r["seq_status"]) == {0}


def test_get_multi_batch_sizes_one_million_groups_fast(lib):
    """O(windows) driver calls were the reference's cost (allocate_block.hpp:367-369); here one million mixed groups
    are binned in about a second, and binning is monotone: more memory never means more batches."""
    import time

    rng = np.random.default_rng(5)
    n = 1_000_000
    wn = np.full(n, 2, dtype=np.int32)
    lt = rng.choice(np.array([250, 500, 1000, 2000], dtype=np.int32), size=n, p=[0.2, 0.5, 0.2, 0.1])
    rl = np.repeat(lt, 2).astype(np.int32)
    t0 = time.perf_counter()
    sizes, w2b = _multi_batch(lib, rl, wn)
    dt = time.perf_counter() - t0
    assert w2b.min() >= 0 and len(sizes) >= 1
    assert dt < 3.0, "binning one million groups took %.2f s" % dt
    fewer_mem, _ = _multi_batch(lib, rl[:20000], wn[:10000], quota=0.05)
    more_mem, _ = _multi_batch(lib, rl[:20000], wn[:10000], quota=0.9)
    assert len(more_mem) <= len(fewer_mem)


def test_get_multi_batch_sizes_vs_reference_build(lib):
    """The reference's binning on the same groups (its own footprint formula and one cudaMemGetInfo per group): both
    partition the input; the native batches additionally guarantee the fit (documented difference, engine.cu)."""
    if not oraclelib.have_ref_cudapoa():
        pytest.skip("reference build not present on this box")
    ws = generate_windows(300, start_index=51000, length_choices=[250, 500, 1000, 2000],
                          length_probs=[0.2, 0.5, 0.2, 0.1], n_reads_range=(10, 60))
    n_sizes, n_map = _multi_batch(lib, ws.read_lengths, ws.window_n_reads)
    r_sizes, r_map = _multi_batch(parity.ref_lib(), ws.read_lengths, ws.window_n_reads)
    assert n_map.min() >= 0 and r_map.min() >= 0
    assert len(n_sizes) >= 1 and len(r_sizes) >= 1
    # both put the longest windows into the first batch
    assert n_sizes[0]["max_sequence_size"] == int(ws.read_lengths.max())
