"""GPU parity tests (-m gpu): every output limb of the CUDA path, called through the C ABI, is compared
bit-for-bit with the CPU oracle on the same keys, ciphertexts and plaintexts (P1), and decrypted outputs are
compared with the reference's plaintext prover relations (P0: fhe/code_test.go:109-116, fhe/ligero_test.go:164-174,
fhe/ligero.go:554-571)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from helpers import Scenario  # noqa: E402
from oracle import oracle as O  # noqa: E402

T = O.T_DEFAULT


def _first_diff(a, b):
    d = np.argwhere(a != b)
    return "first mismatch at %s of %d mismatches (got %d, want %d)" % (d[0].tolist(), len(d), a[tuple(d[0])], b[tuple(d[0])]) if len(d) else ""


def assert_same(got, want, what=""):
    assert got.shape == want.shape, (got.shape, want.shape)
    assert (got == want).all(), what + ": " + _first_diff(got, want)


_cache = {}


def scenario(logN, rows, cols, k):
    key = (logN, rows, cols, k)
    if key not in _cache:
        _cache[key] = Scenario(logN, rows, cols, k)
    return _cache[key]


# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("logN,k", [(6, 3), (8, 3), (10, 3), (12, 3), (13, 2), (14, 3)])
def test_limb_ntt_intt(logN, k):
    S = scenario(logN, 1 << (logN - 1), 16, k)
    be = S.backend()
    st = be.NewStore(1)
    ct = S.cts[0]
    st.upload(ct)
    from lumenos_b200 import _capi
    L = _capi.lib()
    for limb in range(k):
        _capi.check(L.lmn_limb_ntt(be.ctx, st.h, 0, 1, limb, 1))      # INTT of poly 1
    got = st.download()[0]
    for limb in range(k):
        assert_same(got[1, limb], S.P.intt(ct[1, limb], limb), "INTT limb %d" % limb)
    for limb in range(k):
        _capi.check(L.lmn_limb_ntt(be.ctx, st.h, 0, 1, limb, 0))
    assert_same(st.download()[0], ct, "NTT(INTT(x))")
    be.Close()


@pytest.mark.parametrize("logN,k", [(6, 4), (12, 3)])
def test_operator_level(logN, k):
    S = scenario(logN, 1 << (logN - 1), 16, k)
    P = S.P
    be = S.backend()
    st = be.NewStore(4)
    st.upload(S.cts[:2], 0)
    a, b = S.cts[0], S.cts[1]
    be.Add((st, 0), (st, 1), (st, 2)); assert_same(st.download(2, 1)[0], P.add(a, b), "Add")
    be.Sub((st, 0), (st, 1), (st, 2)); assert_same(st.download(2, 1)[0], P.sub(a, b), "Sub")
    for w in [int(S.roots[0]), int(S.roots[4]), int(S.roots[5]), T - 1, 1, 0, 2**64 - 1]:
        be.Mul((st, 0), w, (st, 3)); assert_same(st.download(3, 1)[0], P.mul_scalar(a, w), "Mul scalar %d" % w)
    be.Mul((st, 0), int(S.roots[4]), (st, 0)); assert_same(st.download(0, 1)[0], P.mul_scalar(a, int(S.roots[4])), "Mul in place")
    st.upload(a[None], 0)
    be.MulNew((st, 0), S.rpt, (st, 2)); assert_same(st.download(2, 1)[0], P.mul_pt(a, S.rpt), "MulNew")
    be.CopyNew((st, 1), (st, 3)); assert_same(st.download(3, 1)[0], b, "CopyNew")
    assert be.MulCounter() == 9
    # Rescale chain
    cur, cur_st = a, st
    scale = 1
    for lvl in range(k - 1, 0, -1):
        nxt = be.NewStore(1, lvl - 1)
        be.Rescale((cur_st, 0), (nxt, 0))
        want, scale = P.rescale(cur, scale)
        assert_same(nxt.download()[0], want, "Rescale from level %d" % lvl)
        cur, cur_st = want, nxt
    assert scale == be.RescaleScale(k - 1, 0)
    be.Close()


@pytest.mark.parametrize("logN,rows,k", [(6, 32, 4), (6, 64, 4), (6, 64, 5), (8, 256, 6), (10, 512, 3),
                                         (12, 4096, 4), (13, 8192, 3), (14, 16384, 4), (14, 8192, 3)])
def test_inner_sum(logN, rows, k):
    S = scenario(logN, rows, 16, k)
    be = S.backend()
    st = be.NewStore(1)
    col = S.P.mul_pt(S.cts[1], S.rpt)
    st.upload(col[None])
    be.InnerSum((st, 0), 1, rows)
    want = S.P.innersum(col, rows, S.gal, S.evks)
    assert_same(st.download()[0], want, "InnerSum rows=%d" % rows)
    be.Close()


def test_inner_sum_missing_key_and_bad_n():
    S = scenario(6, 32, 16, 4)
    from lumenos_b200 import fhe
    be = fhe.ServerBFV(fhe.PrimeField(T, S.roots), fhe.BGVParameters(6, S.P.q, S.P.p, T, S.P.psi_q, S.P.psi_p))
    st = be.NewStore(1); st.upload(S.cts[0][None])
    with pytest.raises(fhe.LumenosError, match="missing Galois key"):
        be.InnerSum((st, 0), 1, 32)
    with pytest.raises(fhe.LumenosError, match="power-of-two"):
        be.InnerSum((st, 0), 1, 24)
    be.Close()


# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("logN,rows,cols,k", [(6, 32, 16, 6), (8, 128, 64, 8), (6, 32, 128, 3), (6, 32, 256, 3),
                                              (6, 32, 1024, 2), (12, 2048, 8, 2)])
def test_encode_bit_exact(logN, rows, cols, k):
    from lumenos_b200 import fhe
    S = scenario(logN, rows, cols, k)
    be = S.backend()
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    enc = fhe.Encode(matrix, rows, 2, be, (zero, 0))
    want = S.P.fhe_encode(S.cts, 2, S.zero, S.roots)
    got = enc.download()
    assert_same(got, want, "Encode %dx%d" % (rows, cols))
    assert be.MulCounter() == O.ntt_counts()[2]
    assert_same(matrix.download(), S.cts, "Encode must not modify the matrix")
    # fhe.NTT on an explicit slice gives the same as Encode on [matrix | zero copies]
    full = np.concatenate([S.cts, np.repeat(S.zero[None], cols, axis=0)])
    inp = be.NewStore(2 * cols).upload(full)
    out = fhe.NTT(inp, 2 * cols, be)
    assert_same(out.download(), want, "NTT")
    if k >= 6:  # P0: decrypt == core.Encode (fhe/code_test.go:109-116)
        exp = np.stack([O.core_encode(S.M[i], 2, S.roots) for i in range(rows)])
        for j in range(2 * cols):
            assert (S.P.decrypt(S.sk, got[j])[:rows] == exp[:, j]).all(), j
    be.Close()


@pytest.mark.parametrize("logN,rows,cols,k,batch", [(6, 32, 16, 6, 10), (6, 64, 16, 5, 64), (8, 256, 64, 8, 64),
                                                    (12, 4096, 8, 3, 6), (12, 2048, 8, 6, 5), (14, 16384, 8, 3, 64)])
def test_commit_prove_bit_exact(logN, rows, cols, k, batch):
    """TestLigeroE2E shape of flow (fhe/ligero_test.go:70-176): Commit -> Prove, every output ciphertext compared"""
    from lumenos_b200 import fhe
    S = scenario(logN, rows, cols, k)
    P = S.P
    be = S.backend()
    be.SetBatch(batch)
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    prover = committer.Commit(matrix, be, (zero, 0))
    rng = np.random.default_rng(7)
    idx = rng.integers(0, 2 * cols, size=committer.Queries).astype(np.uint32)
    idx[5] = idx[2]                                                  # duplicates are legal (ligero.go:638-644)
    proof = prover.Prove(S.rpt, S.bpt, idx, be)
    enc_want = P.fhe_encode(S.cts, 2, S.zero, S.roots)
    matR, matZ, qc = proof.MatR.download(), proof.MatZ.download(), proof.QueriedCols.download()
    for j in range(cols):
        wr, sr = P.inner_sum_eval_column(S.cts[j], S.rpt, rows, S.gal, S.evks)
        wz, sz = P.inner_sum_eval_column(S.cts[j], S.bpt, rows, S.gal, S.evks)
        assert_same(matR[j], wr, "MatR[%d]" % j)
        assert_same(matZ[j], wz, "MatZ[%d]" % j)
        assert sr == proof.MatR.scale and sz == proof.MatZ.scale
        # P0: slot 0 == sum_i r_i M_ij , sum_i b_i M_ij (fhe/ligero_test.go:164-174)
        assert int(P.decrypt(S.sk, matR[j], sr)[0]) == sum((int(a) % T) * int(b) for a, b in zip(S.r, S.M[:, j])) % T
        assert int(P.decrypt(S.sk, matZ[j], sz)[0]) == sum(int(a) * int(b) for a, b in zip(S.b, S.M[:, j])) % T
    seen = {}
    for i, ix in enumerate(idx[:40]):
        if int(ix) not in seen:
            seen[int(ix)] = P.rescale_to_level1(enc_want[ix])
        want, sc = seen[int(ix)]
        assert_same(qc[i], want, "QueriedCols[%d] (column %d)" % (i, ix))
        assert sc == proof.QueriedCols.scale
    assert_same(prover.EncodedMatrix.download(), enc_want, "EncodedMatrix")
    be.Close()


@pytest.mark.parametrize("logN,k,n_ct", [(6, 4, 7), (12, 3, 5), (14, 12, 24)])
def test_limb_table_upload_download_roundtrip(logN, k, n_ct):
    """lmn_store_upload_limbs / lmn_store_download_limbs: the host holds one separately allocated slice per limb (Lattigo's
    Coeffs[limb], core/utils.go:12-33); the batched calls must move exactly the bytes the per-limb and the contiguous calls move,
    for ranges inside the store, for chunks smaller and larger than the 32 MiB staging buffer (24 cts at D = 72 MiB), back to back"""
    from lumenos_b200 import fhe
    S = scenario(6, 32, 16, 4) if logN == 6 else None
    if S is None:
        from lumenos_b200 import hostutil as H
        q, p = H.generate_moduli(logN, k)
        params = fhe.BGVParameters(logN, q, p, T, [H.psi_of(m, logN) for m in q], [H.psi_of(m, logN) for m in p])
        be = fhe.ServerBFV(fhe.PrimeField(T, H.field_roots_forward(16, T)), params)
    else:
        be = S.backend()
    N, L = 1 << logN, k - 1
    rng = np.random.default_rng(logN)
    st = be.NewStore(n_ct + 2, L)
    base = rng.integers(0, 2**60, size=(n_ct + 2, 2, k, N), dtype=np.uint64)
    st.upload(base)
    # separately allocated limbs, not adjacent in memory (a spacer between consecutive allocations)
    limbs, spacers = [], []
    for _ in range(n_ct * 2 * k):
        limbs.append(rng.integers(0, 2**50, size=N, dtype=np.uint64))
        spacers.append(np.empty(257, np.uint8))
    st.upload_limbs(limbs, first=1, count=n_ct)
    want = base.copy()
    want[1:1 + n_ct] = np.stack(limbs).reshape(n_ct, 2, k, N)
    assert_same(st.download(), want, "store after the batched limb upload")
    assert_same(st.download_limb(1, 1, k - 1), limbs[2 * k - 1], "one limb through the per-limb call")
    # a kernel queued after the upload sees the data (stream order, no host synchronisation in between)
    st.upload_limbs(limbs[::-1], first=1, count=n_ct)
    be.Add((st, 1), (st, 1), (st, 0))                                     # st[0] = st[1] + st[1]: values below 2^50, so no reduction happens
    out = [np.zeros(N, np.uint64) for _ in range((n_ct + 2) * 2 * k)]
    st.download_limbs(out)
    want[1:1 + n_ct] = np.stack(limbs[::-1]).reshape(n_ct, 2, k, N)
    got = np.stack(out).reshape(n_ct + 2, 2, k, N)
    assert_same(got[1:], want[1:], "batched limb download")
    assert_same(got[0], want[1] * np.uint64(2), "a kernel queued behind the upload read the uploaded limbs")
    with pytest.raises(fhe.LumenosError, match="out of bounds"):
        st.upload_limbs(limbs, first=3, count=n_ct)
    with pytest.raises(ValueError):
        st.upload_limbs(limbs[:-1], first=1, count=n_ct)
    be.Close()


def test_trim_releases_the_pools_and_results_do_not_change():
    """lmn_ctx_trim: a context that has committed and proved holds its batch workspaces and the serialised leaves; trimming returns
    them to the driver (free device memory grows by what the call reports), and the next Commit / Prove regrows them and produces
    the same bytes"""
    import torch
    from lumenos_b200 import fhe
    logN, rows, cols, k = 12, 2048, 16, 3
    S = scenario(logN, rows, cols, k)
    be = S.backend()
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    idx = np.arange(committer.Queries, dtype=np.uint32) % (2 * cols)

    def run():
        prover = committer.Commit(matrix, be, (zero, 0))
        proof = prover.Prove(S.rpt, S.bpt, idx, be)
        return proof.Root, proof.MatR.download(), proof.MatZ.download(), proof.QueriedCols.download()

    first = run()
    be.Synchronize()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    released = be.Trim()
    free1 = torch.cuda.mem_get_info()[0]
    leaf_bytes = 2 * cols * 2 * 2 * (1 << logN) * 8                        # the serialised leaves alone
    assert released >= leaf_bytes, (released, leaf_bytes)
    assert free1 - free0 >= released * 0.8, (free0, free1, released)        # the driver rounds allocations to its own granularity
    assert be.Trim() == 0                                                   # nothing left to give back
    second = run()
    assert first[0] == second[0]
    for a, b, what in zip(first[1:], second[1:], ("MatR", "MatZ", "QueriedCols")):
        assert_same(b, a, what + " after trim")
    be.Close()


def test_ring_switch_bit_exact():
    from lumenos_b200 import fhe
    logN, logNs, rows, cols, k = 8, 6, 128, 16, 4
    S = scenario(logN, rows, cols, k)
    P = S.P
    Ps = O.Params(logNs, q=P.q[:1], p=P.p)
    sks = Ps.keygen_sk(9)
    evk = P.keygen_ringswitch(S.sk, sks, logNs, 11)
    be = S.backend()
    be.SetRingSwitchServer(fhe.RingSwitchServer(be, evk, logNs))
    matrix = be.NewStore(cols).upload(S.cts)
    out = fhe.matrixInnerSumEval(matrix, S.rpt, rows, be)
    got = out.download()
    assert got.shape == (cols, 2, 1, 1 << logNs)
    for j in range(cols):
        lvl1, _ = P.inner_sum_eval_column(S.cts[j], S.rpt, rows, S.gal, S.evks)
        assert_same(got[j], P.ringswitch(lvl1, evk, logNs), "ring switch column %d" % j)
    be.Close()


def test_store_roundtrip_and_errors():
    from lumenos_b200 import fhe
    S = scenario(6, 32, 16, 4)
    be = S.backend()
    st = be.NewStore(3)
    st.upload(S.cts[:3])
    assert_same(st.download(), S.cts[:3], "bulk roundtrip")
    st.upload_limb(2, 1, 3, S.cts[0][0, 1])
    assert_same(st.download_limb(2, 1, 3), S.cts[0][0, 1], "limb roundtrip")
    with pytest.raises(fhe.LumenosError):
        st.upload(S.cts[:4])
    with pytest.raises(fhe.LumenosError):
        be.NewStore(1, level=7)
    lvl1 = be.NewStore(1, 1)
    with pytest.raises(fhe.LumenosError, match="share one level"):
        be.Add((st, 0), (st, 1), (lvl1, 0))
    be.Close()


LEAF_FORMATS = [("lattigo-like, odd header", b"\x01MetaData{json-ish}\x00\x07", None, None), ("bare", b"", b"", b""),
                ("unaligned everywhere", b"abc", b"\x05", b"\x09\x08\x07"), ("aligned", b"A" * 48, None, None)]


@pytest.mark.parametrize("logN,rows,cols,k,batch", [(6, 32, 16, 6, 10), (8, 256, 8, 4, 64), (12, 2048, 8, 3, 3), (6, 32, 8, 3, 64)])
@pytest.mark.parametrize("fmt_id", range(len(LEAF_FORMATS)))
def test_commit_merkle_bit_exact(logN, rows, cols, k, batch, fmt_id):
    """Commit incl. the "Merkle tree built" span (fhe/ligero.go:105-116, 126-183; core/tree.go): leaves, leaf bytes, every leaf hash,
    root and paths against the oracle; paths accepted by the reference's VerifyMerklePath rule; query stage from the leaves."""
    from lumenos_b200 import fhe
    S = scenario(logN, rows, cols, k)
    P = S.P
    be = S.backend()
    be.SetBatch(batch)
    _, hdr, ph, lh = LEAF_FORMATS[fmt_id]
    N = 1 << logN
    lf = fhe.LeafFormat.with_prefixes(hdr, 1, N) if ph is None else fhe.LeafFormat(hdr, ph, lh)
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    n = 2 * cols
    enc_want = P.fhe_encode(S.cts, 2, S.zero, S.roots)
    leaves_want = [P.rescale_to_level1(enc_want[j])[0] for j in range(n)]
    got_leaves = prover.Leaves.download()
    bufs = []
    for j in range(n):
        assert_same(got_leaves[j], leaves_want[j], "leaf %d" % j)
        bufs.append(O.leaf_bytes(leaves_want[j], lf.header, lf.polyHeader, lf.limbHeader))
    for j in (0, n - 1):
        assert fhe.LeafBytes(be, prover.Leaves, j, lf) == bufs[j], "serialised leaf %d" % j
    tree = O.MerkleTree(bufs)
    for j in range(n):
        assert prover.Tree.LeafHash(j) == tree.leaf_hash(j), "leaf hash %d" % j
    assert prover.Root == tree.root()
    rng = np.random.default_rng(11)
    idx = rng.integers(0, n, size=committer.Queries).astype(np.uint32)
    idx[5] = idx[2]
    proof = prover.Prove(S.rpt, S.bpt, idx, be)
    qc = proof.QueriedCols.download()
    assert proof.Root == tree.root() and len(proof.MerklePaths) == len(idx)
    for i, ix in enumerate(idx[:48]):
        assert_same(qc[i], leaves_want[ix], "QueriedCols[%d]" % i)
        assert proof.MerklePaths[i] == tree.path(int(ix))
        assert fhe.VerifyMerklePath(bufs[ix], proof.MerklePaths[i], proof.Root, int(ix))
    with pytest.raises(IndexError):
        prover.Tree.GetMerklePath(n)
    be.Close()


@pytest.mark.parametrize("n_leaves", [1, 2, 3, 5, 33])
def test_merkle_odd_leaf_counts(n_leaves):
    """core.NewTree pairs an unpaired last node with itself (tree.go:122-130): trees over 1, 3, 5, 33 leaves"""
    from lumenos_b200 import fhe
    S = scenario(6, 32, 16, 3)
    be = S.backend()
    rng = np.random.default_rng(n_leaves)
    cts = rng.integers(0, 2**55, size=(n_leaves, 2, 2, 64), dtype=np.uint64)
    leaves = be.NewStore(n_leaves, 1).upload(cts)
    lf = fhe.LeafFormat.with_prefixes(b"hdr", 1, 64)
    t = fhe.MerkleTree(be, leaves, lf)
    bufs = [O.leaf_bytes(cts[j], lf.header, lf.polyHeader, lf.limbHeader) for j in range(n_leaves)]
    want = O.MerkleTree(bufs)
    assert t.MerkleRoot() == want.root()
    for j in range(n_leaves):
        assert t.GetMerklePath(j) == want.path(j)
        assert fhe.VerifyMerklePath(bufs[j], t.GetMerklePath(j), t.MerkleRoot(), j)
    be.Close()


@pytest.mark.parametrize("logN,logNs,k", [(10, 8, 3), (12, 10, 3), (13, 10, 4), (14, 10, 3)])
def test_ring_switch_large_rings_bit_exact(logN, logNs, k):
    """RingSwitchNew (fhe/ring_switch.go:106-113) on level-1 ciphertexts of the rings of configs A-D -> LogN 10 (config E): the
    level-0 key switch runs through the block pipeline (log N >= 12), then INTT / decimate / small NTT"""
    from lumenos_b200 import fhe
    P = O.Params(logN, k=k)
    Ps = O.Params(logNs, q=P.q[:1], p=P.p)
    sk, sks = P.keygen_sk(5), Ps.keygen_sk(9)
    evk = P.keygen_ringswitch(sk, sks, logNs, 11)
    roots = O.field_roots_forward(32)
    params = fhe.BGVParameters(logN, P.q, P.p, P.t, P.psi_q, P.psi_p)
    be = fhe.ServerBFV(fhe.PrimeField(P.t, roots), params)
    be.SetRingSwitchServer(fhe.RingSwitchServer(be, evk, logNs))
    rng = np.random.default_rng(logN)
    n_ct = 5
    cts = np.stack([np.stack([np.stack([rng.integers(0, int(P.q[l]), size=P.N, dtype=np.uint64) for l in range(2)]) for _ in range(2)]) for _ in range(n_ct)])
    be.SetBatch(3)
    st = be.NewStore(n_ct, 1).upload(cts)
    got = be.RingSwitch().RingSwitchNew(st).download()
    assert got.shape == (n_ct, 2, 1, 1 << logNs)
    for j in range(n_ct):
        assert_same(got[j], P.ringswitch(cts[j], evk, logNs), "ring switch ct %d" % j)
    be.Close()


@pytest.mark.parametrize("logN,rows,cols,k,batch", [(6, 32, 16, 4, 6), (12, 2048, 16, 3, 8)])
def test_host_overlapped_entry_points_match(logN, rows, cols, k, batch):
    """lmn_encode_host (limb-wise upload behind the Encode kernels) and lmn_inner_product2_host (batch-wise download behind the
    inner-product kernels) give the bits of lmn_store_upload + lmn_encode and lmn_inner_product2 + lmn_store_download"""
    import ctypes as C
    from lumenos_b200 import _capi
    S = scenario(logN, rows, cols, k)
    be = S.backend()
    be.SetBatch(batch)
    lib = _capi.lib()
    vp = C.c_void_p
    n = 2 * cols
    zero = be.NewStore(1).upload(S.zero[None])
    ref_m = be.NewStore(cols).upload(S.cts)
    ref_e = be.NewStore(n)
    plan = be._plan(cols, 2)
    _capi.check(lib.lmn_encode(be.ctx, plan, ref_m.h, zero.h, 0, ref_e.h))
    r_r, r_z = be.NewStore(cols, 1), be.NewStore(cols, 1)
    _capi.check(lib.lmn_inner_product2(be.ctx, ref_m.h, S.rpt.ctypes.data_as(vp), S.bpt.ctypes.data_as(vp), rows, r_r.h, r_z.h))
    want_e, want_r, want_z = ref_e.download(), r_r.download(), r_z.download()

    def pinned(nbytes):
        p = vp()
        _capi.check(lib.lmn_host_alloc(nbytes, C.byref(p)))
        return p
    h_m = pinned(S.cts.nbytes)
    C.memmove(h_m, S.cts.ctypes.data, S.cts.nbytes)
    h_r, h_z = pinned(want_r.nbytes), pinned(want_z.nbytes)
    m2, e2 = be.NewStore(cols), be.NewStore(n)
    o_r, o_z = be.NewStore(cols, 1), be.NewStore(cols, 1)
    for rep in range(2):
        C.memset(h_r, 0, want_r.nbytes); C.memset(h_z, 0, want_z.nbytes)
        _capi.check(lib.lmn_encode_host(be.ctx, plan, h_m, m2.h, zero.h, 0, e2.h))
        _capi.check(lib.lmn_inner_product2_host(be.ctx, m2.h, S.rpt.ctypes.data_as(vp), S.bpt.ctypes.data_as(vp), rows, o_r.h, o_z.h, h_r, h_z))
        be.Synchronize()
        assert_same(m2.download(), S.cts, "uploaded matrix")
        assert_same(e2.download(), want_e, "EncodedMatrix via lmn_encode_host")
        got_r = np.frombuffer((C.c_uint8 * want_r.nbytes).from_address(h_r.value), dtype=np.uint64).reshape(want_r.shape)
        got_z = np.frombuffer((C.c_uint8 * want_z.nbytes).from_address(h_z.value), dtype=np.uint64).reshape(want_z.shape)
        assert_same(got_r, want_r, "MatR in the host buffer")
        assert_same(got_z, want_z, "MatZ in the host buffer")
        assert_same(o_r.download(), want_r, "MatR store")
    for p in (h_m, h_r, h_z):
        lib.lmn_host_free(p)
    be.Close()


@pytest.mark.parametrize("fmt_id", range(len(LEAF_FORMATS)))
@pytest.mark.parametrize("logN,rows,cols,k", [(6, 32, 16, 6), (8, 256, 8, 4)])
def test_proof_marshal_bit_exact(logN, rows, cols, k, fmt_id):
    """EncryptedProof.MarshalBinary (fhe/ligero.go:646-704) of a real Commit -> Prove: every byte against the oracle's WriteTo,
    for aligned / unaligned / empty framings (a ciphertext stream need not be a multiple of 8 bytes), and read back by ReadFrom."""
    from lumenos_b200 import fhe
    S = scenario(logN, rows, cols, k)
    be = S.backend()
    _, hdr, ph, lh = LEAF_FORMATS[fmt_id]
    N = 1 << logN
    lf = fhe.LeafFormat.with_prefixes(hdr, 1, N) if ph is None else fhe.LeafFormat(hdr, ph, lh)
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    idx = np.random.default_rng(3).integers(0, 2 * cols, size=committer.Queries).astype(np.uint32)
    proof = prover.Prove(S.rpt, S.bpt, idx, be)
    got = proof.MarshalBinary(be, lf)
    matR, matZ, qc = proof.MatR.download(), proof.MatZ.download(), proof.QueriedCols.download()
    fmt = (lf.header, lf.polyHeader, lf.limbHeader)
    want = O.proof_bytes(rows, cols, 2, committer.Queries, matR, matZ, qc, fmt, paths=proof.MerklePaths, root=proof.Root)
    assert len(got) == len(want)
    assert got == want, "first differing byte %d" % next(i for i in range(len(want)) if got[i] != want[i])
    md, r2, z2, q2, p2, root2 = fhe.EncryptedProof.UnmarshalBinary(got, be.params, lf)
    assert (md.Rows, md.Cols, md.RhoInv, md.Queries) == (rows, cols, 2, committer.Queries)
    assert_same(r2, matR, "MatR"); assert_same(z2, matZ, "MatZ"); assert_same(q2, qc, "QueriedCols")
    assert p2 == proof.MerklePaths and root2 == proof.Root
    # a caller-provided buffer that is too small is an error, not a truncated proof
    with pytest.raises(fhe.LumenosError):
        proof.MarshalBinary(be, lf, out=np.empty(len(want) - 1, dtype=np.uint8))
    out = np.zeros(len(want) + 5, dtype=np.uint8)
    assert proof.MarshalBinary(be, lf, out=out) == len(want) and out[:len(want)].tobytes() == want and not out[len(want):].any()
    be.Close()


def test_store_marshal_ranges_and_chunks():
    """lmn_store_marshal: any ciphertext range (the multi-GPU host marshals each rank's columns into its own byte range), streams
    longer than one 64 MiB staging chunk (two alternating buffers), odd stream lengths, and the ring-switched (level-0) shape."""
    import ctypes as C
    from lumenos_b200 import fhe, _capi
    S = scenario(6, 32, 16, 3)
    be = S.backend()
    rng = np.random.default_rng(21)

    def marshal(store, first, count, lf):
        f = lf.c_struct()
        n = C.c_size_t()
        fhe.check(_capi.lib().lmn_store_marshal(be.ctx, store.h, first, count, C.byref(f), None, 0, C.byref(n)))
        buf = np.empty(n.value, dtype=np.uint8)
        fhe.check(_capi.lib().lmn_store_marshal(be.ctx, store.h, first, count, C.byref(f), C.c_void_p(buf.ctypes.data), buf.nbytes, C.byref(n)))
        return buf.tobytes()
    cts = rng.integers(0, 2**58, size=(37, 2, 2, 64), dtype=np.uint64)
    st = be.NewStore(37, 1).upload(cts)
    for lf in (fhe.LeafFormat(b"abc", b"\x05", b"\x09\x08\x07"), fhe.LeafFormat(), fhe.LeafFormat.with_prefixes(b"m" * 13, 1, 64)):
        ser = [O.leaf_bytes(c, lf.header, lf.polyHeader, lf.limbHeader) for c in cts]
        assert marshal(st, 0, 37, lf) == b"".join(ser)
        assert marshal(st, 5, 9, lf) == b"".join(ser[5:14])
        assert marshal(st, 36, 1, lf) == ser[36]
        assert marshal(st, 3, 0, lf) == b""
    with pytest.raises(fhe.LumenosError):
        marshal(st, 30, 8, fhe.LeafFormat())
    # three staging chunks: 2200 level-1 ciphertexts of N = 4096 -> 144 MB + odd framing
    if True:
        S12 = scenario(12, 2048, 8, 3)
        be12 = S12.backend()
        big = rng.integers(0, 2**58, size=(2200, 2, 2, 4096), dtype=np.uint64)
        sb = be12.NewStore(2200, 1).upload(big)
        lf = fhe.LeafFormat(b"abc", b"\x05", b"\x09\x08\x07")
        f = lf.c_struct()
        n = C.c_size_t()
        fhe.check(_capi.lib().lmn_store_marshal(be12.ctx, sb.h, 0, 2200, C.byref(f), None, 0, C.byref(n)))
        buf = np.empty(n.value, dtype=np.uint8)
        fhe.check(_capi.lib().lmn_store_marshal(be12.ctx, sb.h, 0, 2200, C.byref(f), C.c_void_p(buf.ctypes.data), buf.nbytes, C.byref(n)))
        L = n.value // 2200
        for j in (0, 1, 511, 512, 1023, 1024, 1025, 2047, 2048, 2199):    # around the 64 MiB chunk boundaries and the ends
            assert buf[j * L:(j + 1) * L].tobytes() == O.leaf_bytes(big[j], lf.header, lf.polyHeader, lf.limbHeader), "ciphertext %d" % j
        be12.Close()
    be.Close()


@pytest.mark.parametrize("logN,rows,cols,k", [(8, 256, 16, 6), (8, 128, 16, 6)])
def test_ligero_e2e_marshal_decrypt_verify(logN, rows, cols, k):
    """TestLigeroE2E (fhe/ligero_test.go:70-176) end to end: Commit -> Prove -> MarshalBinary -> UnmarshalBinary -> Decrypt -> Verify,
    then MatR / MatZ against the plaintext prover.  The acceptance test is the reference's own Proof.Verify (fhe/ligero.go:516-571)
    restated on the decrypted values: Merkle path of every queried leaf, <col, r> == Encode(MatR)[idx], <col, b> == Encode(MatZ)[idx],
    <MatZ, a> == P(z)."""
    from lumenos_b200 import fhe
    S = scenario(logN, rows, cols, k)
    P = S.P
    be = S.backend()
    lf = fhe.LeafFormat.with_prefixes(b"\x01meta", 1, 1 << logN)
    matrix = be.NewStore(cols).upload(S.cts)
    zero = be.NewStore(1).upload(S.zero[None])
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    idx = np.random.default_rng(17).integers(0, 2 * cols, size=committer.Queries).astype(np.uint32)
    proof = prover.Prove(S.rpt, S.bpt, idx, be)
    wire = proof.MarshalBinary(be, lf)                                                   # server -> client (ligero_test.go:118-126)
    md, matR, matZ, qc, paths, root = fhe.EncryptedProof.UnmarshalBinary(wire, be.params, lf)
    assert (md.Rows, md.Cols, md.RhoInv, md.Queries) == (rows, cols, 2, committer.Queries)
    # EncryptedProof.Decrypt (ligero.go:381-502): MatR / MatZ -> slot 0, queried columns -> their `rows` slots
    sc = proof.MatR.scale
    dR = np.array([int(P.decrypt(S.sk, matR[j], sc)[0]) for j in range(cols)], dtype=np.uint64)
    dZ = np.array([int(P.decrypt(S.sk, matZ[j], sc)[0]) for j in range(cols)], dtype=np.uint64)
    # Proof.Verify (ligero.go:516-571)
    encR, encZ = O.core_encode(dR, 2, S.roots), O.core_encode(dZ, 2, S.roots)
    r = [int(x) % T for x in S.r]
    b = [int(x) for x in S.b]
    a = [pow(S.z, i, T) for i in range(cols)]
    assert pow(S.z, cols, T) == (b[1] if rows > 1 else 1)                               # "zPow is not equal to powA"
    for i, ix in enumerate(idx[:64]):
        leaf = O.leaf_bytes(qc[i], lf.header, lf.polyHeader, lf.limbHeader)               # QueriedCols[i].Ct.WriteTo
        assert fhe.VerifyMerklePath(leaf, paths[i], root, int(ix)), "merkle path of column %d" % ix
        col = [int(v) for v in P.decrypt(S.sk, qc[i], proof.QueriedCols.scale)[:rows]]
        assert sum(c * x for c, x in zip(col, r)) % T == int(encR[ix]), "well-formedness R check failed for column %d" % ix
        assert sum(c * x for c, x in zip(col, b)) % T == int(encZ[ix]), "well-formedness B check failed for column %d" % ix
    value = sum(int(S.M[i, j]) * b[i] % T * a[j] for i in range(rows) for j in range(cols)) % T   # P(z) of the committed polynomial
    assert sum(int(x) * y for x, y in zip(dZ, a)) % T == value, "claimed value does not match the evaluation of the committed polynomial"
    # LigeroProveReference (ligero_test.go:164-174)
    for j in range(cols):
        assert int(dR[j]) == sum(x * int(m) for x, m in zip(r, S.M[:, j])) % T
        assert int(dZ[j]) == sum(x * int(m) for x, m in zip(b, S.M[:, j])) % T
    # a tampered wire byte inside a queried column breaks its Merkle path
    bad = bytearray(wire)
    off_q = 11 + 2 * cols * (len(wire) - 11 - 32 - 32 * len(paths) * len(paths[0])) // (2 * cols + committer.Queries)
    bad[off_q + len(lf.header) + len(lf.polyHeader) + len(lf.limbHeader) + 3] ^= 1
    _, _, _, qc_bad, _, _ = fhe.EncryptedProof.UnmarshalBinary(bytes(bad), be.params, lf)
    assert not fhe.VerifyMerklePath(O.leaf_bytes(qc_bad[0], lf.header, lf.polyHeader, lf.limbHeader), paths[0], root, int(idx[0]))
    be.Close()


@pytest.mark.parametrize("logN,rows,cols,k", [(8, 256, 16, 5), (12, 4096, 8, 4)])
def test_two_host_threads_one_context(logN, rows, cols, k):
    """The reference issues InnerProduct(Matrix, r) and InnerProduct(Matrix, b) from two goroutines on shallow copies of one
    evaluator (fhe/ligero.go:231-242, fhe/bfv.go:56-58).  include/lumenos_b200.h promises the same for one context and distinct
    stores, and a per-thread lmn_last_error: two host threads call the ABI concurrently (ctypes releases the GIL), repeatedly,
    while a third keeps provoking errors; every result must equal the oracle's and every thread must read its own error text."""
    import ctypes as C
    import threading
    from lumenos_b200 import _capi
    S = scenario(logN, rows, cols, k)
    P, lib = S.P, _capi.lib()
    be = S.backend()
    be.SetBatch(6)
    matrix = be.NewStore(cols).upload(S.cts)
    vp = C.c_void_p
    want = {"r": np.stack([P.inner_sum_eval_column(S.cts[j], S.rpt, rows, S.gal, S.evks)[0] for j in range(cols)]),
            "b": np.stack([P.inner_sum_eval_column(S.cts[j], S.bpt, rows, S.gal, S.evks)[0] for j in range(cols)])}
    errors, stop = [], threading.Event()

    def worker(tag, pt):
        try:
            for _ in range(3):
                out = be.NewStore(cols, 1)
                _capi.check(lib.lmn_inner_product(be.ctx, matrix.h, pt.ctypes.data_as(vp), rows, out.h))
                got = out.download()
                if not np.array_equal(got, want[tag]):
                    errors.append("%s: result differs from the oracle under concurrency" % tag)
                # this thread's own failure: an out-of-range download; the message must be this one, not the other thread's
                rc = lib.lmn_store_download(out.h, cols, 1, got.ctypes.data_as(vp))
                msg = lib.lmn_last_error().decode()
                if rc == 0 or "download range" not in msg:
                    errors.append("%s: expected its own download error, got rc=%d %r" % (tag, rc, msg))
                out.free()
        except Exception as e:  # noqa: BLE001
            errors.append("%s: %r" % (tag, e))

    def heckler():
        bad = be.NewStore(1, 0)
        while not stop.is_set():
            rc = lib.lmn_ct_rescale(be.ctx, bad.h, 0, bad.h, 0)          # level-0 input: always invalid
            msg = lib.lmn_last_error().decode()
            if rc == 0 or "Rescale output" not in msg:
                errors.append("heckler: rc=%d %r" % (rc, msg))
                break
        bad.free()

    ts = [threading.Thread(target=worker, args=("r", S.rpt)), threading.Thread(target=worker, args=("b", S.bpt)), threading.Thread(target=heckler)]
    for t in ts:
        t.start()
    ts[0].join(); ts[1].join()
    stop.set(); ts[2].join()
    assert not errors, errors
    be.Close()


def test_ring_switch_key_with_base2_digits_is_rejected():
    """fhe/ring_switch.go:47-55 asks Lattigo for BaseTwoDecomposition = 13; if Lattigo honoured it the key would carry base-2
    sub-digits (evk.Value[d] of length > 1).  The ABI takes the digit shape and refuses such a key instead of misreading it."""
    from lumenos_b200 import fhe
    S = scenario(8, 128, 16, 4)
    be = S.backend()
    P = S.P
    good = np.zeros((P.beta, 2, P.k + P.np, P.N), np.uint64)
    with pytest.raises(fhe.LumenosError, match="base-2"):
        fhe.RingSwitchServer(be, np.zeros((P.beta, 2) + good.shape[1:], np.uint64), 6)
    with pytest.raises(fhe.LumenosError, match="RNS digits"):
        fhe.RingSwitchServer(be, good[:1], 6)
    fhe.RingSwitchServer(be, good, 6)
    be.Close()


def test_merkle_maximum_leaf_count():
    """65536 leaves = the largest codeword length lmn_plan_create admits: the leaf serialiser folds the leaf index into grid.x (a
    grid.y of 65536 does not launch), the SHA-256 producer / consumer pairs cover 2048 CTAs; root and sampled paths against hashlib."""
    import hashlib
    from lumenos_b200 import fhe
    S = scenario(6, 32, 16, 3)
    be = S.backend()
    n = 65536
    rng = np.random.default_rng(9)
    cts = rng.integers(0, 2**55, size=(n, 2, 2, 64), dtype=np.uint64)
    leaves = be.NewStore(n, 1).upload(cts)
    lf = fhe.LeafFormat.with_prefixes(b"\x01abc", 1, 64)
    t = fhe.MerkleTree(be, leaves, lf)
    le = lambda a: a.astype("<u8").tobytes()  # noqa: E731
    buf = lambda j: lf.header + b"".join(lf.polyHeader + b"".join(lf.limbHeader + le(cts[j, p, l]) for l in range(2)) for p in range(2))  # noqa: E731
    lv = [hashlib.sha256(buf(j)).digest() for j in range(n)]
    for j in (0, 1, 4097, n - 1):
        assert t.LeafHash(j) == lv[j]
    while len(lv) > 1:
        lv = [hashlib.sha256(lv[i] + lv[i + 1]).digest() for i in range(0, len(lv), 2)]
    assert t.MerkleRoot() == lv[0]
    for j in (0, 12345, n - 1):
        assert fhe.VerifyMerklePath(buf(j), t.GetMerklePath(j), t.MerkleRoot(), j)
    be.Close()
