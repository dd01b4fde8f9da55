"""Pins the CPU oracle (oracle/cudapoa_oracle.cpp) against the reference's own known-answer vectors:
cudapoa/tests/Test_CudapoaNW.cu:93-175, Test_CudapoaTopSort.cu:42-52, Test_CudapoaAddAlignment.cu:120-219,
Test_CudapoaGenerateConsensus.cu:89-154 -- replayed through the raw-array entry points with the same harness
quirks (zero-filled buffers, weights stored at slot = source id, edge lists in end->start direction)."""
import ctypes

import numpy as np
import pytest

import oraclelib

E = 50  # CUDAPOA_MAX_NODE_EDGES / CUDAPOA_MAX_NODE_ALIGNMENTS
CAP = 64


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _edges(outgoing):
    """BasicGraph::get_edges (cudapoa/tests/basic_graph.hpp:66-82)."""
    in_e = np.zeros(CAP * E, np.int16)
    out_e = np.zeros(CAP * E, np.int16)
    in_c = np.zeros(CAP, np.uint16)
    out_c = np.zeros(CAP, np.uint16)
    for i, outs in enumerate(outgoing):
        out_c[i] = len(outs)
        for j, o in enumerate(outs):
            in_e[o * E + in_c[o]] = i
            in_c[o] += 1
            out_e[i * E + j] = o
    return in_e, in_c, out_e, out_c


def _out_lists(out_e, out_c, n):
    return [[int(out_e[i * E + j]) for j in range(out_c[i])] for i in range(n)]


NW_CASES = [
    ("3,2,1,0", "3,2,1,0", "AAAA", [0, 1, 2, 3], [[1], [2], [3], []], "AATA"),
    ("-1,3,2,1,0", "4,3,2,1,0", "ATCG", [0, 1, 2, 3], [[1], [2], [3], []], "ATCGA"),
    ("3,2,1,0", "3,2,1,0", "AACGC", [0, 4, 1, 2, 3], [[1, 4], [2], [3], [], [2]], "ATCG"),
    ("4,3,2,1,0", "1,-1,-1,-1,0", "ATTGA", [0, 1, 2, 3, 4], [[1], [2], [3], [4], []], "AA"),
    ("4,3,6,5,0", "4,3,2,1,0", "ATGTACA", [0, 5, 1, 6, 2, 3, 4], [[1, 5], [2], [3], [4], [], [6], [3]], "ACTTA"),
]


@pytest.mark.parametrize("case", NW_CASES)
def test_full_nw_known_answers(case):
    ans_graph, ans_read, nodes, order, outgoing, read = case
    lib = oraclelib.oracle()
    n = len(nodes)
    nodes_a = np.frombuffer(nodes.encode(), np.uint8).copy()
    sorted_a = np.array(order, np.int16)
    pos = np.zeros(CAP, np.int16)
    for p, node in enumerate(order):
        pos[node] = p
    in_e, in_c, out_e, out_c = _edges(outgoing)
    read_a = np.frombuffer(read.encode(), np.uint8).copy()
    ag = np.zeros(CAP, np.int16)
    ar = np.zeros(CAP, np.int16)
    width = ((len(read) + 1 + 4 + 3) // 4) * 4
    lib.oracle_nw_full.restype = ctypes.c_int32
    length = lib.oracle_nw_full(_ptr(nodes_a), _ptr(sorted_a), _ptr(pos), n, _ptr(in_c), _ptr(in_e), _ptr(out_c),
                                _ptr(read_a), len(read), width, -8, -6, 8, _ptr(ag), _ptr(ar))
    assert ",".join(str(int(x)) for x in ag[:length]) == ans_graph
    assert ",".join(str(int(x)) for x in ar[:length]) == ans_read


TOPSORT_CASES = [
    ("4-5-0-2-3-1", [[], [], [3], [1], [0, 1], [0, 2]]),
    ("0-1-2-3-4-5", [[1, 3], [2, 3], [3, 4, 5], [4, 5], [5], []]),
    ("6-4-7-5-0-2-3-1", [[], [], [3], [1], [0, 1, 7], [0, 2], [4], [5]]),
]


@pytest.mark.parametrize("case", TOPSORT_CASES)
def test_kahn_topsort_known_answers(case):
    answer, outgoing = case
    lib = oraclelib.oracle()
    n = len(outgoing)
    in_e, in_c, out_e, out_c = _edges(outgoing)
    sorted_a = np.zeros(CAP, np.int16)
    pos = np.zeros(CAP, np.int16)
    lib.oracle_topsort_kahn(_ptr(sorted_a), _ptr(pos), n, _ptr(in_c), _ptr(out_e), _ptr(out_c))
    assert "-".join(str(int(x)) for x in sorted_a[:n]) == answer
    for p in range(n):
        assert pos[sorted_a[p]] == p


ADD_CASES = [
    ([[], [0], [1], [2, 4], [1]], "AAAA", [[], [0], [1], [2]], [1, 1, 1, 1], "AATA", [0, 1, 2, 3], [0, 1, 2, 3]),
    ([[], [0], [1], [2], [3]], "ATCG", [[], [0], [1], [2]], [1, 1, 1, 1], "ATCGA", [0, 1, 2, 3, -1], [0, 1, 2, 3, 4]),
    ([[], [0], [1, 4, 5], [2], [0], [0]], "AACGC", [[], [0], [1, 4], [2], [0]], [2, 1, 2, 2, 1], "ATCG", [0, 4, 2, 3],
     [0, 1, 2, 3]),
    ([[], [0], [1], [2], [3, 0]], "ATTGA", [[], [0], [1], [2], [3]], [1, 1, 1, 1, 1], "AA", [0, 1, 2, 3, 4],
     [0, -1, -1, -1, 1]),
    ([[], [0], [1], [2, 6, 7], [3], [0], [5], [5]], "ATGTACA", [[], [0], [1], [2, 6], [3], [0], [5]],
     [2, 1, 1, 2, 2, 1, 1], "ACTTA", [0, 5, 6, 3, 4], [0, 1, 2, 3, 4]),
]


@pytest.mark.parametrize("case", ADD_CASES)
def test_add_alignment_known_answers(case):
    answer, nodes, outgoing, cov, read, aln_graph, aln_read = case
    lib = oraclelib.oracle()
    n = len(nodes)
    nodes_a = np.zeros(CAP, np.uint8)
    nodes_a[:n] = np.frombuffer(nodes.encode(), np.uint8)
    in_e, in_c, out_e, out_c = _edges(outgoing)
    aligned = np.zeros(CAP * E, np.int16)
    aligned_c = np.zeros(CAP, np.uint16)
    in_w = np.zeros(CAP * E, np.uint16)
    cov_a = np.zeros(CAP, np.uint16)
    cov_a[:n] = cov
    read_a = np.frombuffer(read.encode(), np.uint8).copy()
    ag = np.array(aln_graph, np.int16)
    ar = np.array(aln_read, np.int16)
    weights = np.zeros(len(read), np.int8)  # BasicAlignment never fills base_weights_ (Test_CudapoaAddAlignment.cu:33-42)
    count = ctypes.c_int32(n)
    rc = lib.oracle_add_alignment(_ptr(nodes_a), ctypes.byref(count), _ptr(aligned), _ptr(aligned_c), _ptr(in_e),
                                  _ptr(in_c), _ptr(out_e), _ptr(out_c), _ptr(in_w), len(aln_graph), _ptr(ag),
                                  _ptr(read_a), _ptr(ar), _ptr(cov_a), _ptr(weights), 3072)
    assert rc == 0
    assert _out_lists(out_e, out_c, count.value) == answer


CONSENSUS_CASES = [
    ("ATAA", "AAAAT", [0, 1, 2, 4, 3], [[], [], [4], [], [2]], [[1], [2, 4], [3], [], [3]], [2, 2, 1, 2, 1],
     [[5], [4, 3], [2], [], [1]]),
    ("AGCTA", "ATCGA", [0, 1, 2, 3, 4], [[], [], [], [], []], [[1], [2], [3], [4], []], [1, 1, 1, 1, 1],
     [[4], [3], [2], [1], []]),
    ("GCCA", "AACGCT", [0, 1, 4, 5, 2, 3], [[], [4, 5], [], [], [1, 5], [1, 4]], [[1, 4, 5], [2], [3], [], [2], [2]],
     [3, 1, 3, 3, 1, 1], [[7, 6, 5], [4], [3], [], [2], [1]]),
    ("AGTTA", "ATTGA", [0, 1, 2, 3, 4], [[], [], [], [], []], [[1, 4], [2], [3], [4], []], [2, 1, 1, 1, 2],
     [[5, 4], [3], [2], [1], []]),
    ("ATTCA", "ATGTACAT", [0, 1, 5, 2, 6, 7, 3, 4], [[], [5], [6, 7], [], [], [1], [2, 7], [2, 6]],
     [[1, 5], [2], [3], [4], [], [6, 7], [3], [3]], [3, 1, 1, 3, 3, 2, 1, 1],
     [[9, 8], [7], [6], [5], [], [4, 3], [2], [1]]),
]


@pytest.mark.parametrize("case", CONSENSUS_CASES)
def test_consensus_known_answers(case):
    answer, nodes, order, alignments, outgoing, cov, out_w = case
    lib = oraclelib.oracle()
    n = len(nodes)
    nodes_a = np.frombuffer(nodes.encode(), np.uint8).copy()
    sorted_a = np.array(order, np.int16)
    pos = np.zeros(CAP, np.int16)
    for p, node in enumerate(order):
        pos[node] = p
    in_e, in_c, out_e, out_c = _edges(outgoing)
    aligned = np.zeros(CAP * E, np.int16)
    aligned_c = np.zeros(CAP, np.uint16)
    for i, al in enumerate(alignments):
        for j, a in enumerate(al):
            aligned[i * E + j] = a
            aligned_c[i] += 1
    # BasicGenerateConsensus::get_incoming_edge_w stores the weight at slot = SOURCE NODE ID
    # (Test_CudapoaGenerateConsensus.cu:56-67); the other slots stay zero
    in_w = np.zeros(CAP * E, np.uint16)
    for i, outs in enumerate(outgoing):
        for j, to in enumerate(outs):
            in_w[to * E + i] = out_w[i][j]
    cov_a = np.zeros(CAP, np.uint16)
    cov_a[:n] = cov
    cons = np.zeros(2048 + 8, np.uint8)
    cov_out = np.zeros(2048 + 8, np.uint16)
    rc = lib.oracle_consensus_from_graph(_ptr(nodes_a), n, _ptr(sorted_a), _ptr(pos), _ptr(in_e), _ptr(in_c),
                                         _ptr(out_e), _ptr(out_c), _ptr(in_w), _ptr(cov_a), _ptr(aligned),
                                         _ptr(aligned_c), 2048, _ptr(cons), _ptr(cov_out))
    assert rc == 0
    got = bytes(cons[:list(cons).index(0)]).decode()
    assert got == answer  # the reference's expected strings are the reversed path (end -> start)
