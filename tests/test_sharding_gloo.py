"""CPU: the N > 1 plumbing of bench.py (window ownership, barrier, max/sum over ranks) with world_size 2 over gloo.
No device work: windows are independent, so this is everything the ranks ever exchange."""
import os
import socket
import subprocess
import sys

import pytest

from claragenomicsanalysis_b200.pipeline import shard_bounds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import sys
sys.path.insert(0, %r)
from claragenomicsanalysis_b200.distributed import Ranks
from claragenomicsanalysis_b200.pipeline import shard_bounds
r = Ranks(backend="gloo")
lo, hi = shard_bounds(1001, r.world, r.rank)
r.barrier()
total = r.sum(hi - lo)
slowest = r.max(10.0 + r.rank)
bounds = r.gather_ints([lo, hi])
if r.rank == 0:
    print("RESULT", int(total), slowest, bounds)
r.close()
"""


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("n,k", [(10, 3), (7, 8), (1000000, 8), (0, 4), (5, 1)])
def test_shard_bounds_partition(n, k):
    covered = []
    sizes = []
    for s in range(k):
        lo, hi = shard_bounds(n, k, s)
        assert 0 <= lo <= hi <= n
        covered.extend(range(lo, hi) if n < 100 else [lo, hi])
        sizes.append(hi - lo)
    assert sum(sizes) == n
    assert max(sizes) - min(sizes) <= 1
    if n < 100:
        assert covered == list(range(n))
    else:
        for s in range(1, k):
            assert covered[2 * s] == covered[2 * s - 1]  # contiguous


def test_two_ranks_over_gloo():
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER % ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=300) for p in procs]
    for p, (out, err) in zip(procs, outs):
        assert p.returncode == 0, err[-2000:]
    line = [l for l in outs[0][0].splitlines() if l.startswith("RESULT")][0]
    assert line == "RESULT 1001 11.0 [0, 501, 501, 1001]"
