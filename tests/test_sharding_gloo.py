"""World-size-2 (and 4) CPU tests of the multi-GPU host logic over gloo: the partitions cover the work exactly once
and the queried-column all-to-all rebuilds whole ciphertexts bit-for-bit."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from lumenos_b200 import sharding as S
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n, k, N, nq = 64, 3, 128, 37
g = torch.Generator().manual_seed(5)
enc = torch.randint(0, 2**62, (n, 2, k, N), generator=g, dtype=torch.int64)        # same on every rank
qidx = np.random.default_rng(3).integers(0, n, size=nq); qidx[4] = qidx[1]
lo, hi = S.coefficient_range(N, rank, world)
full = S.exchange_queried_torch(enc[..., lo:hi].contiguous(), qidx, rank, world)
mine = S.queries_of(nq, rank, world)
assert full.shape == (len(mine), 2, k, N)
assert torch.equal(full, enc[torch.as_tensor(qidx[mine])]), "rank %%d: reassembled ciphertexts differ" %% rank
# whole encoded matrix: coefficient slices -> column blocks (Commit's leaf span)
for chunks in (1, 3, 8):
    blk = S.exchange_encoded_torch(enc[..., lo:hi].contiguous(), rank, world, chunks=chunks)
    a, b = S.column_range(n, rank, world)
    assert blk.shape == (b - a, 2, k, N) and torch.equal(blk, enc[a:b]), "rank %%d: column block differs" %% rank
# partitions
cols = 32
owned = torch.zeros(cols, dtype=torch.int64); a, b = S.column_range(cols, rank, world); owned[a:b] += 1
dist.all_reduce(owned); assert bool((owned == 1).all())
q = torch.zeros(nq, dtype=torch.int64); q[mine] += 1
dist.all_reduce(q); assert bool((q == 1).all())
print("rank", rank, "ok")
'''


@pytest.mark.parametrize("world", [2, 4])
def test_exchange_and_partitions(world, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    port = 29600 + world
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == world


def test_partition_errors():
    from lumenos_b200 import sharding as S
    with pytest.raises(ValueError):
        S.column_range(10, 0, 4)
    with pytest.raises(ValueError):
        S.coefficient_range(64, 0, 4)
    assert S.queries_of(7, 1, 3) == [1, 4]
    assert S.column_range(4096, 7, 8) == (3584, 4096)
