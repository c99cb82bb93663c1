"""Multi-GPU equivalence (-m gpu; a case is skipped when the box has fewer devices than its world size): the sharded path
(coefficient-sliced Encode, column-sharded inner products, all-to-all of queried / encoded slices, gathered Merkle tree, sharded
wire sections) produces byte-identical outputs to the oracle's single-process evaluation, at world sizes 2 / 4 / 8, on a small
ring (whole-limb kernels) and on a logN-12 ring with rows == N (the block pipeline of the benchmark, InnerSum's row-swap step)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
import ctypes as C
import numpy as np, torch, torch.distributed as dist
from helpers import Scenario
from lumenos_b200 import fhe, sharding, _capi
from oracle import oracle as O
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
S = Scenario(logN=int(os.environ["LMN_T_LOGN"]), rows=int(os.environ["LMN_T_ROWS"]), cols=int(os.environ["LMN_T_COLS"]), k=int(os.environ["LMN_T_K"]))
P, lib = S.P, _capi.lib()
be = S.backend(device=local, stream=torch.cuda.current_stream(dev).cuda_stream)
want_enc = P.fhe_encode(S.cts, 2, S.zero, S.roots)
# Encode on coefficient slices
lo, hi = sharding.coefficient_range(P.N, rank, world)
m_sl = torch.from_numpy(S.cts[..., lo:hi].astype(np.int64)).contiguous().to(dev)
z_sl = torch.from_numpy(S.zero[None][..., lo:hi].astype(np.int64)).contiguous().to(dev)
e_sl = torch.empty((2 * S.cols, 2, P.k, hi - lo), dtype=torch.int64, device=dev)
vp = C.c_void_p
_capi.check(lib.lmn_encode_lanes(be.ctx, be._plan(S.cols, 2), vp(m_sl.data_ptr()), vp(z_sl.data_ptr()), vp(e_sl.data_ptr()), (hi - lo).bit_length() - 1))
be.Synchronize()
assert np.array_equal(e_sl.cpu().numpy().astype(np.uint64), want_enc[..., lo:hi]), "sliced Encode differs"
# the library's communicator (NCCL behind the C ABI), bootstrapped through torch.distributed
sharding.init_comm(be, rank, world)
assert lib.lmn_comm_rank(be.ctx) == rank and lib.lmn_comm_world(be.ctx) == world
# query columns via the library's all-to-all (lmn_exchange_queried), cross-checked against the torch.distributed twin
qidx = np.random.default_rng(3).integers(0, 2 * S.cols, size=37); qidx[4] = qidx[1]
mine = sharding.queries_of(len(qidx), rank, world)
st = sharding.exchange_queried(be, e_sl, qidx)
full = sharding.exchange_queried_torch(e_sl, qidx, rank, world)
assert np.array_equal(st.download()[:len(mine)], full.cpu().numpy().astype(np.uint64)), "library and torch all-to-all disagree"
out = be.NewStore(max(1, len(mine)), 1)
idx = np.arange(len(mine), dtype=np.uint32)
_capi.check(lib.lmn_query_columns(be.ctx, st.h, idx.ctypes.data_as(vp), len(mine), out.h))
got = out.download()
for j, qi in enumerate(mine):
    assert np.array_equal(got[j], P.rescale_to_level1(want_enc[qidx[qi]])[0]), "queried column %%d differs" %% qi
# inner products on the column shard
a, b = sharding.column_range(S.cols, rank, world)
mat = be.NewStore(b - a).upload(S.cts[a:b])
mr, mz = be.NewStore(b - a, 1), be.NewStore(b - a, 1)
_capi.check(lib.lmn_inner_product2(be.ctx, mat.h, S.rpt.ctypes.data_as(vp), S.bpt.ctypes.data_as(vp), S.rows, mr.h, mz.h))
gr, gz = mr.download(), mz.download()
for j in range(a, b):
    assert np.array_equal(gr[j - a], P.inner_sum_eval_column(S.cts[j], S.rpt, S.rows, S.gal, S.evks)[0])
    assert np.array_equal(gz[j - a], P.inner_sum_eval_column(S.cts[j], S.bpt, S.rows, S.gal, S.evks)[0])
# config E on the column shard (fhe/ligero.go:336-342: the ring switch sits inside the per-column worker loop, so it shards with the
# columns): every rank loads the ring-switch key and switches its own MatR / MatZ block; no exchange is involved
logNs = 10 if P.logN >= 12 else 6
Ps = O.Params(logNs, q=P.q[:1], p=P.p)
rs_evk = P.keygen_ringswitch(S.sk, Ps.keygen_sk(9), logNs, 11)
be.SetRingSwitchServer(fhe.RingSwitchServer(be, rs_evk, logNs))
for store, lvl1 in ((mr, gr), (mz, gz)):
    small = be.RingSwitch().RingSwitchNew(store).download()
    assert small.shape == (b - a, 2, 1, 1 << logNs)
    for j in range(b - a):
        assert np.array_equal(small[j], P.ringswitch(lvl1[j], rs_evk, logNs)), "ring-switched column %%d differs on rank %%d" %% (a + j, rank)
# wire format on the column shards: every rank marshals its own MatR / MatZ block, the blocks concatenated in rank order are the
# MatR / MatZ sections of EncryptedProof.WriteTo (fhe/ligero.go:665-679)
from oracle import oracle as O
wf = fhe.LeafFormat(b"abc", b"\x05", b"\x09\x08\x07")      # odd lengths: streams are not multiples of 8 bytes
parts_r, parts_z = [None] * world, [None] * world
dist.all_gather_object(parts_r, sharding.marshal_columns(be, mr, wf))
dist.all_gather_object(parts_z, sharding.marshal_columns(be, mz, wf))
want_r = b"".join(O.leaf_bytes(P.inner_sum_eval_column(S.cts[j], S.rpt, S.rows, S.gal, S.evks)[0], wf.header, wf.polyHeader, wf.limbHeader) for j in range(S.cols))
want_z = b"".join(O.leaf_bytes(P.inner_sum_eval_column(S.cts[j], S.bpt, S.rows, S.gal, S.evks)[0], wf.header, wf.polyHeader, wf.limbHeader) for j in range(S.cols))
assert b"".join(parts_r) == want_r and b"".join(parts_z) == want_z, "sharded MatR / MatZ wire sections differ from the oracle"
# Commit's Merkle span, sharded: slices -> column blocks (lmn_exchange_encoded), local leaf rescale + leaf hashes, gathered tree
n = 2 * S.cols
a2, b2 = sharding.column_range(n, rank, world)
for batch in (64, 3):                       # the batch size bounds the exchange chunk: one chunk, and several with a remainder
    be.SetBatch(batch)
    blk = sharding.exchange_encoded(be, e_sl)
    assert np.array_equal(blk.download(), want_enc[a2:b2]), "column block of the encoded matrix differs (batch %%d)" %% batch
    blk_t = sharding.exchange_encoded_torch(e_sl, rank, world, chunks=2)
    assert np.array_equal(blk_t.cpu().numpy().astype(np.uint64), want_enc[a2:b2])
    # the pipelined form: leaves straight from the slices == exchange + lmn_commit_leaves
    leaves = sharding.commit_leaves_from_slices(be, e_sl)
    leaves2 = be.NewStore(b2 - a2, 1)
    _capi.check(lib.lmn_commit_leaves(be.ctx, blk.h, leaves2.h))
    want_leaves = np.stack([P.rescale_to_level1(want_enc[j])[0] for j in range(a2, b2)])
    assert np.array_equal(leaves.download(), want_leaves) and np.array_equal(leaves2.download(), want_leaves), "leaves differ (batch %%d)" %% batch
    # and back: column block -> coefficient slices reproduces this rank's slice of the encoded matrix
    back = torch.empty_like(e_sl)
    sharding.to_slices(be, blk, n, back)
    be.Synchronize()
    assert torch.equal(back, e_sl), "block -> slices does not invert slices -> block (batch %%d)" %% batch
lf = fhe.LeafFormat.with_prefixes(b"hdr-bytes", 1, P.N)
tree = sharding.commit_tree(be, leaves, lf)
bufs = [O.leaf_bytes(P.rescale_to_level1(want_enc[j])[0], lf.header, lf.polyHeader, lf.limbHeader) for j in range(n)]
want_tree = O.MerkleTree(bufs)
assert tree.MerkleRoot() == want_tree.root(), "sharded Merkle root differs from the oracle"
for j in (0, a2, b2 - 1, n - 1):
    assert tree.GetMerklePath(j) == want_tree.path(j)
# the matrix itself: uploaded once as a column block, its coefficient slices derived over NVLink feed the same Encode
a, b = sharding.column_range(S.cols, rank, world)
mat_blk = be.NewStore(b - a).upload(S.cts[a:b])
m_sl2 = torch.empty_like(m_sl)
sharding.to_slices(be, mat_blk, S.cols, m_sl2)
be.Synchronize()
assert torch.equal(m_sl2, m_sl), "matrix slices derived from the column blocks differ from the host-sliced ones"
# query stage from the local leaves
own = [int(q) for q in qidx if a2 <= q < b2]
if own:
    sel = be.NewStore(len(own), 1)
    li = np.array([q - a2 for q in own], dtype=np.uint32)
    _capi.check(lib.lmn_store_gather(be.ctx, leaves.h, li.ctypes.data_as(vp), len(own), sel.h))
    gs = sel.download()
    for t, q in enumerate(own):
        assert np.array_equal(gs[t], P.rescale_to_level1(want_enc[q])[0])
be.Close()
dist.barrier()
print("rank", rank, "ok")
dist.destroy_process_group()
'''


CASES = [(2, 8, 256, 16, 5), (2, 12, 4096, 16, 5), (4, 12, 4096, 16, 5), (8, 12, 4096, 16, 4), (8, 8, 128, 16, 5)]


@pytest.mark.parametrize("world,logN,rows,cols,k", CASES)
def test_multi_gpu_equivalence(tmp_path, world, logN, rows, cols, k):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, ROOT))
    env = dict(os.environ, LMN_T_LOGN=str(logN), LMN_T_ROWS=str(rows), LMN_T_COLS=str(cols), LMN_T_K=str(k))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(29655 + world), str(script)],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == world
