"""CPU: the oracle against outputs of the UNMODIFIED reference cudapoa (built for sm_100a with its own flags and run
on a B200 by tests/golden/make_reference_golden.py).  Together with test_oracle_golden.py this pins the oracle."""
import pytest

import oraclelib
import parity
from cases import CASES, case_windows


def _golden(name):
    g = parity.load_golden("reference_%s.json" % name)
    if g is None:
        pytest.skip("golden fixture for %s not generated" % name)
    return g


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_outputs(name):
    g = _golden(name)
    case = CASES[name]
    if case["engine"]["banded"] and oraclelib.rcp_table() is None:
        pytest.skip("MUFU.RCP table missing: banded band starts cannot be reproduced bit-exactly on the CPU")
    ws = case_windows(name)
    assert len(ws) == g["n_windows"]
    output = case["engine"]["output"]
    o = parity.run_oracle(ws, want_graphs=True, **case["engine"])
    assert [int(s) for s in o["status"]] == g["status"]
    if output == "consensus":
        assert o["consensus"] == g["consensus"]
        assert [list(c) for c in o["coverage"]] == g["coverage"]
    else:
        assert o["msa"] == g["msa"]
    assert parity.digests(o, output) == g["digest"]
    # graphs: node labels of the first windows, all edge lists stored, node and edge counts of every window
    for w, labels in enumerate(g["graph_nodes"]):
        assert o["graphs"][w][0].decode("latin1") == labels
    for w, edges in enumerate(g["graph_edges"]):
        assert [list(e) for e in o["graphs"][w][1]] == edges
    assert [len(gr[0]) for gr in o["graphs"]] == g["graph_node_count"]
    assert [len(gr[1]) for gr in o["graphs"]] == g["graph_edge_count"]
    assert o["rcp_misses"] == 0 or not case["engine"]["banded"]


def test_score_and_size_type_selection():
    """create_batch picks int16/int32 like cudapoa_limits.hpp:28-53 ([probe] SURVEY.md 3.1: int32 scores from
    L >= 1490 with -8/-6/8, int32 sizes only from banded L >= 8192)."""
    from claragenomicsanalysis_b200.windows import generate_windows

    ws = generate_windows(1, template_length=40, n_reads=2)
    for max_seq, banded, bits in ((1489, True, (16, 16)), (1490, True, (32, 16)), (8191, True, (32, 16)),
                                  (8192, True, (32, 32)), (8192, False, (32, 16)), (10923, False, (32, 32))):
        p = oraclelib.make_params(max_seq=max_seq, max_reads=2, band=256, banded=banded)
        r = oraclelib.OracleRun(ws, p)
        assert (r.score_bits, r.size_bits) == bits, (max_seq, banded)


def test_stale_score_memory_never_matters():
    """The reference never clears its score buffer and banded alignment reads cells it has not written
    (cudapoa_nw_banded.cuh:77-80, :224-248).  Two different fill patterns must give identical outputs."""
    ws = case_windows("banded16_noisy").slice(0, 6)
    eng = CASES["banded16_noisy"]["engine"]
    a = parity.run_oracle(ws, stale_fill=0x00, **eng)
    b = parity.run_oracle(ws, stale_fill=0x7B, **eng)
    c = parity.run_oracle(ws, stale_fill=0xFF, **eng)
    assert parity.compare(a, b, "consensus") == []
    assert parity.compare(a, c, "consensus") == []


def test_spoa_agreement_rate_reported():
    """Agreement with the reference SPOA (kNW 8/-6/-8) is a reported number, not an equality: cudapoa breaks
    ties differently.  Guard only against a collapse."""
    if not oraclelib.have_ref_spoa():
        pytest.skip("oracle/_ref/libspoa_ref.so not built")
    ws = case_windows("full16_consensus")
    o = parity.run_oracle(ws, **CASES["full16_consensus"]["engine"])
    sp = oraclelib.SpoaRun(ws)
    same = sum(o["consensus"][w] == sp.consensus(w) for w in range(len(ws)))
    assert same >= len(ws) // 2
