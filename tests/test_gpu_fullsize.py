"""Full-size parity (-m gpu): the CUDA path at BASELINE.json's shapes (ring degree, chain length and matrix size of
configs B and D) on uniformly random canonical residues, checked
  * bit-for-bit against the oracle on random samples (lanes of Encode, columns of the inner products, query columns),
  * through size-independent properties: Encode is linear per lane, results do not depend on the batch size, and a
    second run reproduces the first."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import oracle as O  # noqa: E402

T = O.T_DEFAULT


def _backend(P, roots):
    from lumenos_b200 import fhe
    import torch
    params = fhe.BGVParameters(P.logN, P.q, P.p, P.t, P.psi_q, P.psi_p)
    # the stores wrap torch tensors, so the library runs on torch's current stream (one ordering for both)
    return fhe.ServerBFV(fhe.PrimeField(P.t, roots), params, stream=torch.cuda.current_stream().cuda_stream)


def _rand_cts(torch, dev, q, count, N, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    t = torch.empty((count, 2, len(q), N), dtype=torch.int64, device=dev)
    for i, m in enumerate(q):
        t[:, :, i, :] = torch.randint(0, int(m), (count, 2, N), generator=g, dtype=torch.int64, device=dev)
    return t


@pytest.mark.parametrize("logN,cols,k", [(12, 2048, 11), (14, 4096, 12)])
def test_encode_full_size_sampled_lanes_and_linearity(logN, cols, k):
    import torch
    from lumenos_b200 import _capi, fhe
    dev = torch.device("cuda", 0)
    P = O.Params(logN, k=k)
    N, n = P.N, 2 * cols
    roots = O.field_roots_forward(n)
    be = _backend(P, roots)
    lib = _capi.lib()
    A = _rand_cts(torch, dev, P.q, cols, N, 1)
    Z = _rand_cts(torch, dev, P.q, 1, N, 2)
    EA = torch.empty((n, 2, k, N), dtype=torch.int64, device=dev)
    sA, sZ = fhe.CiphertextStore(be, cols, k - 1, wrap_ptr=A.data_ptr()), fhe.CiphertextStore(be, 1, k - 1, wrap_ptr=Z.data_ptr())
    sEA = fhe.CiphertextStore(be, n, k - 1, wrap_ptr=EA.data_ptr())
    plan = be._plan(cols, 2)
    _capi.check(lib.lmn_encode(be.ctx, plan, sA.h, sZ.h, 0, sEA.h))
    be.Synchronize()
    # ---- bit-exact against the oracle on random lanes ----
    rng = np.random.default_rng(7)
    nl = 1024
    lanes = np.sort(rng.choice(2 * k * N, size=nl, replace=False))
    limb_of_lane = ((lanes // N) % k).astype(np.int32)
    lt = torch.as_tensor(lanes, device=dev)
    ml = A.view(cols, -1)[:, lt].cpu().numpy().astype(np.uint64)
    zl = Z.view(1, -1)[0, lt].cpu().numpy().astype(np.uint64)
    got = EA.view(n, -1)[:, lt].cpu().numpy().astype(np.uint64)
    want = P.fhe_encode_lanes(ml, 2, zl, limb_of_lane, roots)
    assert np.array_equal(got, want), "Encode differs from the oracle on sampled lanes (%d mismatches)" % int((got != want).sum())
    # ---- a second run reproduces the first ----
    first = EA.view(n, -1)[:, lt].clone()
    torch.cuda.synchronize()   # the clone runs on torch's stream, the library on its own: order them
    _capi.check(lib.lmn_encode(be.ctx, plan, sA.h, sZ.h, 0, sEA.h))
    be.Synchronize()
    assert torch.equal(first, EA.view(n, -1)[:, lt])
    # ---- linearity per lane (config B size only: three encoded matrices must fit) ----
    if logN == 12:
        qv = torch.as_tensor(P.q.astype(np.int64), device=dev).view(1, 1, k, 1)
        B = _rand_cts(torch, dev, P.q, cols, N, 3)
        Z2 = _rand_cts(torch, dev, P.q, 1, N, 4)
        EB, ES = torch.empty_like(EA), torch.empty_like(EA)
        S, ZS = (A + B) % qv, (Z + Z2) % qv
        st = lambda t, c: fhe.CiphertextStore(be, c, k - 1, wrap_ptr=t.data_ptr())  # noqa: E731
        sB, sZ2, sS, sZS, sEB, sES = st(B, cols), st(Z2, 1), st(S, cols), st(ZS, 1), st(EB, n), st(ES, n)
        _capi.check(lib.lmn_encode(be.ctx, plan, sB.h, sZ2.h, 0, sEB.h))
        _capi.check(lib.lmn_encode(be.ctx, plan, sS.h, sZS.h, 0, sES.h))
        be.Synchronize()
        assert torch.equal((EA + EB) % qv, ES), "Encode(A) + Encode(B) != Encode(A + B)"
    be.Close()


@pytest.mark.parametrize("logN,rows,k", [(12, 4096, 11), (13, 8192, 12), (14, 16384, 12)])
def test_prove_full_ring_sampled_columns(logN, rows, k):
    """inner products + query rescale at the full ring size and chain length of configs B / D (columns are independent,
    so 48 columns exercise exactly the per-column work of the full matrix)"""
    import torch
    from lumenos_b200 import _capi, fhe
    dev = torch.device("cuda", 0)
    P = O.Params(logN, k=k)
    N, cols = P.N, 48
    roots = O.field_roots_forward(128)   # any valid field: Prove does not use the Encode twiddles
    be = _backend(P, roots)
    lib = _capi.lib()
    rng = np.random.default_rng(11)
    mods = list(P.q) + list(P.p)
    gal = P.galois_elements_for_innersum(rows)
    evks = []
    for g in gal:
        e = np.empty((P.beta, 2, k + P.np, N), np.uint64)
        for i, m in enumerate(mods):
            e[:, :, i, :] = rng.integers(0, int(m), size=(P.beta, 2, N), dtype=np.uint64)
        evks.append(e)
        be.LoadGaloisKey(g, e)
    M = _rand_cts(torch, dev, P.q, cols, N, 5)
    E = _rand_cts(torch, dev, P.q, 2 * cols, N, 6)           # stands for an encoded matrix
    sM, sE = fhe.CiphertextStore(be, cols, k - 1, wrap_ptr=M.data_ptr()), fhe.CiphertextStore(be, 2 * cols, k - 1, wrap_ptr=E.data_ptr())
    rpt = np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in P.q])
    bpt = np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in P.q])
    vp = C.c_void_p
    outs = {}
    for batch in (64, 10):
        be.SetBatch(batch)
        mr, mz = be.NewStore(cols, 1), be.NewStore(cols, 1)
        _capi.check(lib.lmn_inner_product2(be.ctx, sM.h, rpt.ctypes.data_as(vp), bpt.ctypes.data_as(vp), rows, mr.h, mz.h))
        outs[batch] = (mr.download(), mz.download())
    assert np.array_equal(outs[64][0], outs[10][0]) and np.array_equal(outs[64][1], outs[10][1]), "result depends on the batch size"
    # the kernels overlap asynchronous copies with compute: repeated runs must reproduce the result bit for bit
    be.SetBatch(64)
    for rep in range(4):
        mr, mz = be.NewStore(cols, 1), be.NewStore(cols, 1)
        _capi.check(lib.lmn_inner_product2(be.ctx, sM.h, rpt.ctypes.data_as(vp), bpt.ctypes.data_as(vp), rows, mr.h, mz.h))
        assert np.array_equal(mr.download(), outs[64][0]) and np.array_equal(mz.download(), outs[64][1]), "run %d differs from the first" % rep
    Mh = M.cpu().numpy().astype(np.uint64)
    for j in (0, 17, 47):
        wr, _ = P.inner_sum_eval_column(Mh[j], rpt, rows, gal, evks)
        wz, _ = P.inner_sum_eval_column(Mh[j], bpt, rows, gal, evks)
        assert np.array_equal(outs[64][0][j], wr), "MatR[%d] differs from the oracle" % j
        assert np.array_equal(outs[64][1][j], wz), "MatZ[%d] differs from the oracle" % j
    idx = np.array([5, 90, 5, 33], dtype=np.uint32)
    q = be.NewStore(len(idx), 1)
    _capi.check(lib.lmn_query_columns(be.ctx, sE.h, idx.ctypes.data_as(vp), len(idx), q.h))
    got = q.download()
    Eh = E[torch.as_tensor(idx.astype(np.int64), device=dev)].cpu().numpy().astype(np.uint64)
    for i in range(len(idx)):
        assert np.array_equal(got[i], P.rescale_to_level1(Eh[i])[0]), "queried column %d differs from the oracle" % idx[i]
    assert np.array_equal(got[0], got[2])
    be.Close()


@pytest.mark.parametrize("logN,n_leaves", [(12, 4096), (14, 8192)])
def test_merkle_full_size_properties(logN, n_leaves):
    """Commit's Merkle span at the leaf count and leaf size of configs B / D (8192 x 512 KiB): sampled leaf hashes against hashlib
    over the downloaded ciphertext bytes, the root recomputed on the host from the GPU's leaf hashes (core.NewTree), and sampled
    paths accepted by the reference's VerifyMerklePath rule (core/tree.go:223-268)."""
    import hashlib
    import torch
    from lumenos_b200 import fhe
    dev = torch.device("cuda", 0)
    P = O.Params(logN, k=3)
    N = P.N
    be = _backend(P, O.field_roots_forward(128))
    g = torch.Generator(device=dev).manual_seed(21)
    L = torch.randint(0, 2**55, (n_leaves, 2, 2, N), generator=g, dtype=torch.int64, device=dev)
    leaves = fhe.CiphertextStore(be, n_leaves, 1, wrap_ptr=L.data_ptr())
    lf = fhe.LeafFormat.with_prefixes(b"\x07" * 45, 1, N)           # odd header length: every coefficient lands unaligned
    torch.cuda.synchronize()
    tree = fhe.MerkleTree(be, leaves, lf)
    le = lambda a: a.cpu().numpy().astype("<u8").tobytes()  # noqa: E731
    sample = [0, 1, n_leaves // 2 + 3, n_leaves - 1]
    bufs = {}
    for j in sample:
        bufs[j] = lf.header + b"".join(lf.polyHeader + b"".join(lf.limbHeader + le(L[j, p, l]) for l in range(2)) for p in range(2))
        assert tree.LeafHash(j) == hashlib.sha256(bufs[j]).digest(), "leaf hash %d" % j
    lv = [tree.LeafHash(j) for j in range(n_leaves)]
    while len(lv) > 1:
        lv = [hashlib.sha256(lv[i] + (lv[i + 1] if i + 1 < len(lv) else lv[i])).digest() for i in range(0, len(lv), 2)]
    assert tree.MerkleRoot() == lv[0]
    for j in sample:
        assert fhe.VerifyMerklePath(bufs[j], tree.GetMerklePath(j), tree.MerkleRoot(), j)
    # a second build reproduces the root
    assert fhe.MerkleTree(be, leaves, lf).MerkleRoot() == lv[0]
    be.Close()


@pytest.mark.parametrize("rows,cols,logN,k,p1", [(2048, 1024, 12, 10, 59828798142202325), (4096, 2048, 12, 11, 78852759954010476)])
def test_config_A_B_ligero_e2e_verify(rows, cols, logN, k, p1):
    """BASELINE.json configs[0] = the reference's own test, at its own shape: TestLigeroE2E / TestLigeroPPD (fhe/ligero_test.go:13-19,
    70-176), 2048 x 1024, LogN 12, k = log2(cols) = 10 Q primes (GenerateBGVParamsForNTT), rhoInv 2, 309 queries, REAL keys and
    pk-encryptions of the ChaCha20 witness (P(1) known answer); and configs[1], 4096 x 2048 (rows == N: InnerSum's row-swap step).
    Commit -> Prove -> MarshalBinary -> ReadFrom -> Decrypt -> the checks of Proof.Verify (fhe/ligero.go:516-571) -> MatR / MatZ ==
    LigeroProveReference; two MatR columns also bit-for-bit against the oracle."""
    from helpers import Scenario
    from lumenos_b200 import fhe
    assert O.lib().orc_num_q_primes_for_ntt(cols, T) == k                 # fhe/bfv.go:147-154
    S = Scenario(logN, rows, cols, k)
    P = S.P
    assert int(S.M.astype(object).sum() % T) == p1        # results/baseline/client/bench_{2048x1024,4096x2048}_12.txt:22
    be = S.backend()
    lf = fhe.LeafFormat.with_prefixes(b"\x00" * 48, 1, 1 << logN)
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    assert committer.Queries == 309
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    idx = (np.random.default_rng(5).integers(0, 2**62, size=309) % (2 * cols)).astype(np.uint32)
    proof = prover.Prove(S.rpt, S.bpt, idx, be)
    wire = proof.MarshalBinary(be, lf)
    md, matR, matZ, qc, paths, root = fhe.EncryptedProof.UnmarshalBinary(wire, be.params, lf)
    assert (md.Rows, md.Cols, md.RhoInv, md.Queries) == (rows, cols, 2, 309) and len(paths[0]) == (2 * cols - 1).bit_length() and root == prover.Root
    sc = proof.MatR.scale
    dR = np.array([int(P.decrypt(S.sk, matR[j], sc)[0]) for j in range(cols)], dtype=np.uint64)
    dZ = np.array([int(P.decrypt(S.sk, matZ[j], sc)[0]) for j in range(cols)], dtype=np.uint64)
    Mo = S.M.astype(object)
    r = np.array([int(x) % T for x in S.r], dtype=object)
    b = np.array([int(x) for x in S.b], dtype=object)
    a = np.array([pow(S.z, i, T) for i in range(cols)], dtype=object)
    # LigeroProveReference (ligero_test.go:164-174)
    assert [int(x) for x in dR] == [int(x) % T for x in r.dot(Mo)], "MatR differs from the plaintext prover"
    assert [int(x) for x in dZ] == [int(x) % T for x in b.dot(Mo)], "MatZ differs from the plaintext prover"
    # Proof.Verify
    encR, encZ = O.core_encode(dR, 2, S.roots), O.core_encode(dZ, 2, S.roots)
    for i, ix in enumerate(idx):
        leaf = O.leaf_bytes(qc[i], lf.header, lf.polyHeader, lf.limbHeader)
        assert fhe.VerifyMerklePath(leaf, paths[i], root, int(ix)), "failed to verify merkle path for column %d" % ix
        col = P.decrypt(S.sk, qc[i], proof.QueriedCols.scale)[:rows].astype(object)
        assert int(col.dot(r)) % T == int(encR[ix]), "well-formedness R check failed for column %d" % ix
        assert int(col.dot(b)) % T == int(encZ[ix]), "well-formedness B check failed for column %d" % ix
    value = int(b.dot(Mo).dot(a)) % T
    assert int(dZ.astype(object).dot(a)) % T == value, "claimed value does not match the evaluation of the committed polynomial"
    # P1 on two columns
    for j in (0, cols - 1):
        want, _ = P.inner_sum_eval_column(S.cts[j], S.rpt, rows, S.gal, S.evks)
        assert (matR[j] == want).all(), "MatR[%d] differs from the oracle" % j
    be.Close()
