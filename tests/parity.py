"""Helpers shared by the parity tests: run a WindowSet through any library exposing the cudapoa C ABI and return
plain Python results comparable with the oracle's."""
import hashlib
import json
import os

import numpy as np

from claragenomicsanalysis_b200 import _cabi
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch

import oraclelib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

_ref_lib = None


def ref_lib():
    """The reference cudapoa behind the same C ABI (oracle/_ref, test infrastructure)."""
    global _ref_lib
    if _ref_lib is None:
        _ref_lib = _cabi.CabiLibrary(oraclelib.REF_CUDAPOA_SO, "cudapoa_ref_")
    return _ref_lib


def run_abi(lib, ws, max_seq, max_reads, band=256, banded=True, output="consensus", gap=-8, mismatch=-6, match=8,
            max_mem=2 << 30, want_graphs=False, batch_size_c=None, device_id=0):
    """Run all windows of `ws` (several batch rounds if they do not fit at once).
    Returns dict(status=[...], consensus=[...], coverage=[...], msa=[...], graphs=[...], seq_status=array)."""
    batch = CudaPoaBatch(max_reads, max_seq, max_mem, output_type=output, device_id=device_id, gap_score=gap,
                         mismatch_score=mismatch, match_score=match, cuda_banded_alignment=banded,
                         alignment_band_width=band, batch_size_c=batch_size_c, lib=lib)
    out = dict(status=[], consensus=[], coverage=[], msa=[], graphs=[], seq_status=[])
    start = 0
    n = len(ws)
    try:
        while start < n:
            part = ws.slice(start, n)
            rc, added, seq_status = batch.add_poa_groups_flat(part.bases, part.read_lengths, part.window_n_reads,
                                                              weights=part.weights, want_seq_status=True)
            if added == 0:
                raise RuntimeError("batch accepted no window (status %d)" % rc)
            n_reads_added = int(part.window_n_reads[:added].sum())
            out["seq_status"].extend(int(x) for x in seq_status[:n_reads_added])
            batch.generate_poa()
            if output == "consensus":
                cons, cov, st = batch.get_consensus()
                out["consensus"].extend(cons)
                out["coverage"].extend(cov)
            else:
                msa, st = batch.get_msa()
                out["msa"].extend(msa)
            out["status"].extend(st)
            if want_graphs:
                graphs, _ = batch.get_graph_arrays()
                out["graphs"].extend(graphs)
            out.setdefault("stats", []).append(batch.stats())
            batch.reset()
            start += added
    finally:
        batch.close()
    return out


def run_oracle(ws, max_seq, max_reads, band=256, banded=True, output="consensus", gap=-8, mismatch=-6, match=8,
               want_graphs=False, batch_size=None, **kw):
    p = oraclelib.make_params(max_seq=max_seq, max_reads=max_reads, band=band, banded=banded, output=output, gap=gap,
                              mismatch=mismatch, match=match, batch_size=batch_size, **kw)
    r = oraclelib.OracleRun(ws, p)
    out = dict(status=r.statuses(), consensus=[], coverage=[], msa=[], graphs=[], score_bits=r.score_bits,
               size_bits=r.size_bits, dp_cells=r.dp_cells, rcp_misses=r.rcp_misses)
    for w in range(len(ws)):
        if output == "consensus":
            c, cov = r.consensus(w)
            out["consensus"].append(c)
            out["coverage"].append(cov)
        else:
            out["msa"].append(r.msa(w))
        if want_graphs:
            out["graphs"].append(r.graph(w) if out["status"][w] == 0 else (b"", []))
    return out


def window_digest(result, w, output):
    """Stable per-window digest of status + outputs."""
    h = hashlib.sha256()
    h.update(str(int(result["status"][w])).encode())
    if output == "consensus":
        h.update(result["consensus"][w].encode())
        h.update(np.asarray(result["coverage"][w], dtype=np.uint16).tobytes())
    else:
        for row in result["msa"][w]:
            h.update(row.encode())
            h.update(b"\n")
    return h.hexdigest()[:16]


def digests(result, output):
    return [window_digest(result, w, output) for w in range(len(result["status"]))]


def compare(a, b, output, names=("a", "b")):
    """Indices of windows whose status/outputs differ."""
    bad = []
    for w in range(len(a["status"])):
        same = int(a["status"][w]) == int(b["status"][w])
        if same and output == "consensus":
            same = a["consensus"][w] == b["consensus"][w] and list(a["coverage"][w]) == list(b["coverage"][w])
        elif same:
            same = list(a["msa"][w]) == list(b["msa"][w])
        if not same:
            bad.append(w)
    return bad


def load_golden(name):
    path = os.path.join(GOLDEN_DIR, name)
    if not os.path.exists(path):
        return None
    with open(path) as f:
        return json.load(f)
