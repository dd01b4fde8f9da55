"""End-to-end check of the reference (cudapoa/tests/Test_CudapoaBatchEnd2End.cu:33-85): windows cut from a genome go
through several batches at once, the coverage-trimmed consensuses are concatenated (multi_batch.hpp:176-207) and the
result is compared with the genome and, window by window, with the consensus and coverage of the unmodified reference
build.  The reference's window file (cudapoa/data/sample-windows.txt) is not part of the
snapshot, so the windows are simulated here from a synthetic genome of the same size (33 kb): 30 reads per 500 bp
window at 2 % error, the regime in which the reference's binding test expects consensus == truth
(pygenomeworks/test/test_cudapoa_bindings.py:121-146)."""
import numpy as np
import pytest

from claragenomicsanalysis_b200.pipeline import assembly
from claragenomicsanalysis_b200.windows import _ACGT, WindowSet, _mutate

GENOME_LENGTH = 33458  # length of the reference's sample-golden-value.txt
WINDOW = 500


def simulated_windows(seed=20261104):
    rng = np.random.default_rng(seed)
    genome = rng.integers(0, 4, GENOME_LENGTH, dtype=np.int64)
    wins = []
    for a in range(0, GENOME_LENGTH, WINDOW):
        seg = genome[a:a + WINDOW]
        reads = [bytes(_ACGT[seg]).decode()]  # the backbone is the draft segment itself
        for _ in range(29):
            reads.append(bytes(_ACGT[_mutate(rng, seg, 0.006, 0.006, 0.008)]).decode())
        wins.append(reads)
    return bytes(_ACGT[genome]).decode(), WindowSet.from_lists(wins)


def test_assembly_trims_by_average_coverage():
    # average coverage 3: leading 1,2 and trailing 2 are trimmed
    assert assembly(["ACGTAC"], [[1, 2, 4, 4, 5, 2]]) == "GTA"
    # constant coverage keeps everything; a window reduced to one base contributes nothing (begin < end is strict)
    assert assembly(["ACGT", "TTT"], [[3, 3, 3, 3], [0, 9, 0]]) == "ACGT"
    assert assembly([""], [[]]) == ""


@pytest.mark.gpu
@pytest.mark.parametrize("workers", [2, 4])
def test_simulated_genome_is_reassembled(workers):
    from claragenomicsanalysis_b200.pipeline import WindowPipeline

    genome, ws = simulated_windows()
    pipe = WindowPipeline(30, 640, 1 << 30, devices=(0,), workers_per_device=workers, output_type="consensus",
                          cuda_banded_alignment=True, alignment_band_width=256)
    try:
        pipe.run(ws, 8)  # many small rounds per worker, as the reference's multi-batch driver does
        assert all(s == 0 for s in pipe.statuses())
        cons, cov = zip(*(pipe.consensus(i) for i in range(len(ws))))
    finally:
        pipe.close()
    # Most windows' consensus is their genome segment; a read that ends in an insertion can leave a stray low-coverage
    # base after the last backbone node, which is what the coverage trimming of the assembly removes.
    exact = sum(c == genome[w * WINDOW:(w + 1) * WINDOW] for w, c in enumerate(cons))
    assert exact >= 0.6 * len(ws)
    # ... and window by window it is what the UNMODIFIED reference build computes for these windows (when it travelled)
    import oraclelib
    import parity
    if oraclelib.have_ref_cudapoa():
        r = parity.run_abi(parity.ref_lib(), ws, max_mem=4 << 30, max_seq=640, max_reads=30, band=256, banded=True,
                           output="consensus")
        assert r["status"] == [0] * len(ws)
        assert list(cons) == r["consensus"]
        assert [list(c) for c in cov] == [list(c) for c in r["coverage"]]
    asm = assembly(cons, cov)
    assert len(asm) >= 0.97 * len(genome)
    pos = 0
    for w, c in enumerate(cons):  # the assembly is the in-order concatenation of one substring of every window
        seg = genome[w * WINDOW:(w + 1) * WINDOW]
        n = len(cov[w])
        avg = sum(int(x) for x in cov[w]) // n
        b = next(i for i in range(n) if cov[w][i] >= avg)
        e = next(i for i in range(n - 1, -1, -1) if cov[w][i] >= avg)
        piece = c[b:e + 1]
        assert piece in seg and len(piece) >= 0.9 * len(seg), "window %d" % w
        assert asm[pos:pos + len(piece)] == piece
        pos += len(piece)
    assert pos == len(asm)
