"""GPU: the reference's own pygenomeworks Cython module (genomeworks.cudapoa.CudaPoaBatch, cudapoa.pyx / cuda.pyx
cythonised where they lie under /root/reference by oracle/build_pygenomeworks.py, linked against THIS repository's
headers and libcudapoa_b200.so) driving the native engine.  The assertions are those of the reference's
pygenomeworks/test/test_cudapoa_bindings.py:20-146 and test_cuda_bindings.py.  The module exercises the C++ plugin
interface directly (create_batch, BatchSize's 6-argument constructor, the Batch vtable, DirectedGraph, exception
translation through `except +`): a drift in any exported symbol, struct layout or vtable order fails here."""
import os
import random
import sys
from difflib import SequenceMatcher

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYGW = os.path.join(ROOT, "oracle", "_ref", "pygw")


@pytest.fixture(scope="module")
def gw():
    if not os.path.isdir(os.path.join(PYGW, "genomeworks")):
        pytest.skip("oracle/_ref/pygw not built (needs /root/reference at build time: oracle/build_pygenomeworks.py)")
    sys.path.insert(0, PYGW)
    import genomeworks.cuda as cuda
    import genomeworks.cudapoa as cudapoa

    return cuda, cudapoa


def _mem(cuda):
    device = cuda.cuda_get_device()
    free, total = cuda.cuda_get_mem_info(device)
    return device, min(0.9 * free, 8 << 30)


def test_module_is_linked_against_the_native_library(gw):
    cuda, cudapoa = gw
    with open("/proc/self/maps") as f:
        maps = f.read()
    assert "libcudapoa_b200.so" in maps
    assert cudapoa.status_to_str(0) == "success" and cudapoa.status_to_str(8) == "output_type_unavailable"


@pytest.mark.parametrize("banded", [False, True])
def test_cudapoa_simple_batch(gw, banded):  # test_cudapoa_bindings.py:20-51
    cuda, cudapoa = gw
    device, mem = _mem(cuda)
    batch = cudapoa.CudaPoaBatch(10, 1024, mem, device_id=device, output_type="consensus", cuda_banded_alignment=banded)
    batch.add_poa_group(["ACTGACTG", "ACTTACTG", "ACGGACTG", "ATCGACTG"])
    batch.add_poa_group(["ACTGAC", "ACTTAC", "ACGGAC", "ATCGAC"])
    batch.generate_poa()
    consensus, coverage, status = batch.get_consensus()
    assert len(consensus) == 2 and batch.total_poas == 2
    assert [cudapoa.status_to_str(s) for s in status] == ["success", "success"]
    assert consensus[0] == "ACTGACTG" and len(coverage[0]) == 8


def test_cudapoa_output_types(gw):  # :54-74
    cuda, cudapoa = gw
    device, mem = _mem(cuda)
    with pytest.raises(RuntimeError):
        cudapoa.CudaPoaBatch(10, 1024, mem, device_id=device, output_type="error_input")
    b = cudapoa.CudaPoaBatch(10, 1024, mem, device_id=device, output_type="msa")
    b.add_poa_group(["ACGTACGT", "ACGAACGT"])
    b.generate_poa()
    with pytest.raises(RuntimeError):  # StatusType::output_type_unavailable -> RuntimeError (cudapoa.pyx:224-240)
        b.get_consensus()
    msa, status = b.get_msa()
    assert len(msa) == 1 and len(msa[0]) == 2 and status == [0]


def test_cudapoa_reset_batch(gw):  # :77-91
    cuda, cudapoa = gw
    device, mem = _mem(cuda)
    batch = cudapoa.CudaPoaBatch(10, 1024, mem, device_id=device)
    batch.add_poa_group(["ACTGACTG", "ACTTACTG", "ACGGACTG", "ATCGACTG"])
    batch.generate_poa()
    batch.get_consensus()
    assert batch.total_poas == 1
    batch.reset()
    assert batch.total_poas == 0


def test_cudapoa_graph(gw):  # :94-118 -- networkx DiGraph built from DirectedGraph::get_node_ids / get_edges
    cuda, cudapoa = gw
    device, mem = _mem(cuda)
    batch = cudapoa.CudaPoaBatch(10, 1024, mem, device_id=device)
    batch.add_poa_group(["ACTGACTG", "ACTTACTG", "ACTCACTG"])
    batch.generate_poa()
    batch.get_consensus()
    graphs, status = batch.get_graphs()
    assert len(graphs) == 1
    assert graphs[0].number_of_nodes() == 10 and graphs[0].number_of_edges() == 11
    labels = sorted(d["label"] for _, d in graphs[0].nodes(data=True))
    assert labels == sorted("ACTGACTG" + "TC")


def test_cudapoa_complex_batch(gw):  # :121-146, with a CudaStream from the reference's cuda module
    cuda, cudapoa = gw
    random.seed(2)
    ref = "".join(random.choice("ACGT") for _ in range(500))
    reads = ["".join(r if random.random() > 0.02 else random.choice("ACGT") for r in ref) for _ in range(100)]
    device, mem = _mem(cuda)
    stream = cuda.CudaStream()
    batch = cudapoa.CudaPoaBatch(1000, 1024, mem, stream=stream, device_id=device)
    add_status, seq_status = batch.add_poa_group(reads)
    assert add_status == 0 and list(seq_status) == [0] * 100
    batch.generate_poa()
    consensus, coverage, status = batch.get_consensus()
    assert len(consensus[0]) == len(ref)
    assert SequenceMatcher(None, ref, consensus[0]).ratio() == 1.0


def test_cuda_bindings(gw):  # test_cuda_bindings.py
    cuda, _ = gw
    assert cuda.cuda_get_device_count() >= 1
    cuda.cuda_set_device(0)
    assert cuda.cuda_get_device() == 0
    free, total = cuda.cuda_get_mem_info(0)
    assert 0 < free <= total
    s = cuda.CudaStream()
    s.sync()
