"""Parity cases shared by tests/golden/make_reference_golden.py (run on the GPU box against the reference build) and
the parity tests.  Each case fully determines its windows (generator arguments) and engine parameters."""
from claragenomicsanalysis_b200.windows import generate_windows

# name -> dict(gen=generate_windows kwargs, n=windows, engine=run_abi/run_oracle kwargs)
CASES = {
    # BASELINE.json config 2: 500 bp x 30 reads, banded int16 consensus
    "cfg2_banded16_consensus": dict(
        gen=dict(template_length=500, n_reads=30), n=32,
        engine=dict(max_seq=640, max_reads=30, band=256, banded=True, output="consensus")),
    # config 3 (static band): same windows, MSA output
    "cfg3_banded16_msa": dict(
        gen=dict(template_length=500, n_reads=30), n=12,
        engine=dict(max_seq=640, max_reads=30, band=256, banded=True, output="msa")),
    "banded16_weighted": dict(
        gen=dict(template_length=400, n_reads=16, with_weights=True), n=16, start=1000,
        engine=dict(max_seq=512, max_reads=16, band=256, banded=True, output="consensus")),
    "banded16_band128": dict(
        gen=dict(template_length=500, n_reads=12), n=16, start=2000,
        engine=dict(max_seq=640, max_reads=12, band=128, banded=True, output="consensus")),
    "banded16_short_reads": dict(
        gen=dict(template_length=100, n_reads=10), n=16, start=3000,
        engine=dict(max_seq=256, max_reads=10, band=256, banded=True, output="consensus")),
    "banded16_noisy": dict(  # 25 % errors: bubbles, far predecessors, band-edge traffic
        gen=dict(template_length=600, n_reads=20, read_error=0.25, backbone_error=0.15), n=16, start=4000,
        engine=dict(max_seq=896, max_reads=20, band=256, banded=True, output="consensus")),
    "full16_consensus": dict(
        gen=dict(template_length=300, n_reads=12), n=24, start=5000,
        engine=dict(max_seq=384, max_reads=12, band=256, banded=False, output="consensus")),
    "full16_msa": dict(
        gen=dict(template_length=200, n_reads=10), n=12, start=6000,
        engine=dict(max_seq=256, max_reads=10, band=256, banded=False, output="msa")),
    # config 4 shape scaled down: L >= 1490 switches the reference to int32 scores (cudapoa_limits.hpp:28-37)
    "banded32_consensus": dict(
        gen=dict(template_length=1600, n_reads=8), n=6, start=7000,
        engine=dict(max_seq=2048, max_reads=8, band=256, banded=True, output="consensus")),
    "banded32_msa": dict(
        gen=dict(template_length=1600, n_reads=6), n=4, start=7500,
        engine=dict(max_seq=2048, max_reads=6, band=256, banded=True, output="msa")),
    # config 4 lengths: the packed fill stores these rows relative to a per-row offset (poa_fast.cuh, fast_path_beta)
    "banded32_long": dict(
        gen=dict(template_length=4000, n_reads=10), n=3, start=11000,
        engine=dict(max_seq=5120, max_reads=10, band=256, banded=True, output="consensus")),
    "banded32_long_noisy": dict(  # slow score growth: the offset slope overshoots, scores drift downwards
        gen=dict(template_length=2500, n_reads=8, read_error=0.25, backbone_error=0.15), n=3, start=12000,
        engine=dict(max_seq=3584, max_reads=8, band=256, banded=True, output="consensus")),
    "full32_consensus": dict(
        gen=dict(template_length=1550, n_reads=5), n=4, start=8000,
        engine=dict(max_seq=2048, max_reads=5, band=256, banded=False, output="consensus")),
    # BASELINE.json config 4 at its real shape: 2-5 kb templates x 100 reads, BatchSize(6400, 100): int32 scores in the
    # reference, 99 alignments per window, graphs of several thousand nodes
    "cfg4_banded32_real": dict(
        gen=dict(template_length=(2000, 5000), n_reads=100), n=2, start=20000,
        engine=dict(max_seq=6400, max_reads=100, band=256, banded=True, output="consensus")),
    # config 5B: mixed sizes in ONE batch -- templates 250/500/1000/2000 bp (p = .2/.5/.2/.1), 10..60 reads per window
    "cfg5b_mixed": dict(
        gen=dict(length_choices=[250, 500, 1000, 2000], length_probs=[0.2, 0.5, 0.2, 0.1], n_reads_range=(10, 60)),
        n=64, start=30000,
        engine=dict(max_seq=2560, max_reads=60, band=256, banded=True, output="consensus")),
    # banded max_sequence_size >= 8192: int32 node ids and sizes as well (cudapoa_limits.hpp:39-53), <int32, int32>
    "banded_size32": dict(
        gen=dict(template_length=300, n_reads=5), n=3, start=40000,
        engine=dict(max_seq=8192, max_reads=5, band=256, banded=True, output="consensus")),
    "banded_size32_msa": dict(
        gen=dict(template_length=200, n_reads=4), n=2, start=41000,
        engine=dict(max_seq=8192, max_reads=4, band=256, banded=True, output="msa")),
    # more reads than max_sequences_per_poa and reads longer than max_sequence_size: per-sequence rejections
    "banded16_rejections": dict(
        gen=dict(template_length=(300, 500), n_reads=14), n=12, start=9000, long_read_at=3,
        engine=dict(max_seq=512, max_reads=10, band=256, banded=True, output="consensus")),
}


def case_windows(name):
    c = CASES[name]
    ws = generate_windows(c["n"], start_index=c.get("start", 0), **c["gen"])
    if "long_read_at" in c:
        # splice an over-long read (read 1 + read 2) into every window at the given position
        from claragenomicsanalysis_b200.windows import WindowSet

        wins = []
        for i in range(len(ws)):
            reads = ws.window(i)
            reads.insert(c["long_read_at"], reads[1] + reads[2])
            wins.append(reads)
        ws = WindowSet.from_lists(wins)
    return ws


def structural_windows(n, start=11000, template_length=800, n_reads=12, block=(80, 140), every=3):
    """Windows in which every `every`-th read lost a block of `block` bases a fifth of the way in and gained a random
    block of the same size a fifth before its end: between the two the alignment runs `block` cells off the diagonal,
    more than a 128-wide static band follows.  Used by the adaptive-banding tests (no reference counterpart; the
    yardstick is the full alignment)."""
    import numpy as np

    from claragenomicsanalysis_b200.windows import WindowSet

    base = generate_windows(n, start_index=start, template_length=template_length, n_reads=n_reads)
    rng = np.random.default_rng(start)
    acgt = "ACGT"
    wins = []
    for i in range(len(base)):
        reads = base.window(i)
        for r in range(1, len(reads)):
            if r % every == 0:
                rd = reads[r]
                d = int(rng.integers(block[0], block[1] + 1))
                a = len(rd) // 5
                b = max(a + d + 1, 4 * len(rd) // 5)
                ins = "".join(acgt[int(x)] for x in rng.integers(0, 4, d))
                reads[r] = rd[:a] + rd[a + d:b] + ins + rd[b:]
        wins.append(reads)
    return WindowSet.from_lists(wins)
