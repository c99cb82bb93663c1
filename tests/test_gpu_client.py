"""GPU tests (-m gpu) of the stages either side of the hot path, through the C ABI:
  * the device witness generator against the oracle's core.RandomMatrixRowMajor restatement and the reference's P(1) known answers;
  * the device batch encoder bit for bit against the oracle's bgv.Encoder.Encode;
  * the device public-key encryptor at P0: the oracle's client decrypts every column to the witness, with noise no larger than the
    oracle's own encryptor + 1 bit (Lattigo's sampler stream cannot be reproduced, so ciphertext bits are not the contract);
  * the device decryptor + decoder against the oracle's Decrypt + Decode on the prover's real outputs."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from helpers import Scenario, pk_as_lattigo, sk_as_lattigo  # noqa: E402
from oracle import oracle as O  # noqa: E402

T = O.T_DEFAULT
KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat.json")))
SEED = bytes(range(32))


def _server(S):
    be = S.backend()
    be.LoadPublicKey(pk_as_lattigo(S.P, S.pk))
    return be


def test_witness_generator_matches_oracle_and_known_answers():
    from lumenos_b200 import fhe, hostutil as H
    S = Scenario(6, 32, 8, 3)
    be = S.backend()
    for rows, cols in ((32, 8), (33, 7), (64, 24)):
        M = O.witness_matrix(rows, cols)
        assert np.array_equal(be.WitnessColumns(rows, cols), M.T)
        assert np.array_equal(be.WitnessColumns(rows, cols, firstCol=cols // 2, count=2), M.T[cols // 2:cols // 2 + 2])
    be.Close()
    # the reference's printed P(1) = sum of all entries mod t (results/baseline/client/bench_*.txt:22), at every benchmark shape
    for shape, logN in (("2048x1024", 12), ("4096x2048", 12), ("8192x4096", 13), ("16384x4096", 14)):
        rows, cols = map(int, shape.split("x"))
        q, p = H.generate_moduli(logN, 2)
        params = fhe.BGVParameters(logN, q, p, T, [H.psi_of(m, logN) for m in q], [H.psi_of(m, logN) for m in p])
        be = fhe.ServerBFV(fhe.PrimeField(T, H.field_roots_forward(16, T)), params)
        total = 0
        for c0 in range(0, cols, 1024):
            W = be.WitnessColumns(rows, cols, firstCol=c0, count=min(1024, cols - c0))
            total += int((W & np.uint64(0xFFFFFFFF)).sum(dtype=np.uint64)) + (int((W >> np.uint64(32)).sum(dtype=np.uint64)) << 32)
        assert total % T == KAT["witness_p1"][shape], shape
        be.Close()


@pytest.mark.parametrize("logN,k", [(6, 3), (8, 4), (12, 3), (13, 2), (14, 2)])
def test_encode_plaintext_bit_exact(logN, k):
    S = Scenario(logN, 1 << (logN - 1), 2, k)
    be = S.backend()
    rng = np.random.default_rng(logN)
    for n in (1, 5, S.P.N // 2, S.P.N):
        v = rng.integers(0, 2**63, size=n, dtype=np.uint64) * np.uint64(2) + np.uint64(1)      # unreduced, like SampleUints
        assert np.array_equal(be.Encode(v), S.P.encode(v)), (logN, n)
    be.Close()


@pytest.mark.parametrize("logN,rows,cols,k", [(6, 32, 16, 6), (8, 256, 8, 4), (12, 2048, 8, 4), (12, 4096, 8, 3), (13, 8192, 4, 3), (14, 16384, 4, 3)])
def test_encrypt_columns_decrypts_to_the_witness(logN, rows, cols, k):
    S = Scenario(logN, rows, cols, k)
    P = S.P
    be = _server(S)
    M = O.witness_matrix(rows, cols)
    st = be.EncryptColumns(rows, cols, SEED)                       # witness generated on the device
    cts = st.download()
    ref_noise = max(P.noise_bits(S.sk, P.encrypt(S.pk, None, 77 + j)) for j in range(4))
    for j in range(cols):
        assert np.array_equal(P.decrypt(S.sk, cts[j])[:rows], M[:, j]), "column %d does not decrypt to the witness" % j
    zero = be.EncryptZero(SEED).download()[0]
    assert not P.decrypt(S.sk, zero).any()
    assert P.noise_bits(S.sk, zero) <= ref_noise + 1.0, (P.noise_bits(S.sk, zero), ref_noise)
    # the same columns from host values, in two shards: identical ciphertexts (randomness is keyed by the column index)
    a = be.EncryptColumns(rows, cols, SEED, firstCol=0, count=cols // 2, values=M.T[:cols // 2].copy()).download()
    b = be.EncryptColumns(rows, cols, SEED, firstCol=cols // 2, values=M.T[cols // 2:].copy()).download()
    assert np.array_equal(np.concatenate([a, b]), cts)
    # a different seed gives different ciphertexts of the same plaintexts; distinct columns use distinct randomness
    other = be.EncryptColumns(rows, cols, bytes(32)).download()
    assert not np.array_equal(other[0], cts[0]) and np.array_equal(P.decrypt(S.sk, other[0])[:rows], M[:, 0])
    z2 = be.EncryptColumns(rows, 2, SEED, values=np.zeros((2, rows), np.uint64)).download()
    assert not np.array_equal(z2[0], z2[1])
    be.Close()


def _samples(be, seed, ident, zero_column=False):
    import ctypes as C
    from lumenos_b200 import _capi
    out = np.empty((3, be.params.N), np.int64)
    _capi.check(_capi.lib().lmn_test_encryption_samples(be.ctx, bytes(seed), ident, int(zero_column), out.ctypes.data_as(C.c_void_p)))
    return out


def test_encryption_sampler_distribution_and_its_tie_to_the_ciphertext():
    """The security-relevant half of the public-key encryptor (rlwe.Encryptor: u uniform ternary, e0 / e1 ~ lattigo DefaultXe, sigma 3.2
    bound 19): the polynomials the device draws are dumped (lmn_test_encryption_samples) and checked as distributions, and the dump is
    tied to the real ciphertext by the exact identity c0 + c1 s = u e_pk + e0 + e1 s of an encryption of zero."""
    logN, k, ncol = 12, 3, 8
    S = Scenario(logN, 2048, 4, k)
    P, N = S.P, S.P.N
    be = _server(S)
    D = np.stack([_samples(be, SEED, j) for j in range(ncol)])            # [ncol][3][N]
    u, e0, e1 = D[:, 0].ravel(), D[:, 1].ravel(), D[:, 2].ravel()
    n = u.size                                                            # 32768 samples per polynomial kind
    assert set(np.unique(u)) == {-1, 0, 1}
    for v in (-1, 0, 1):                                                  # sd of a frequency = 0.0026
        assert abs((u == v).mean() - 1 / 3) < 0.012, (v, (u == v).mean())
    for e in (e0, e1):
        assert np.abs(e).max() <= 19 and np.abs(e).max() >= 9             # truncated at 6 sigma, and not a narrow stand-in
        assert abs(e.mean()) < 0.09, e.mean()                             # sd of the mean = 0.018
        assert abs(e.std() - 3.2) < 0.08, e.std()                         # rounding adds 1/12 to the variance: 3.213; sd of the estimate 0.0125
        assert abs((np.abs(e) <= 3).mean() - 0.726) < 0.02                # P(|round(3.2 Z)| <= 3) = P(|Z| < 1.094)
    assert not np.array_equal(D[:, 1], D[:, 2]) and abs(np.corrcoef(e0, e1)[0, 1]) < 0.03      # e0 and e1 are different draws
    assert abs(np.corrcoef(e0[:-1], e0[1:])[0, 1]) < 0.03                 # neighbours of one Box-Muller pair are independent
    assert len({D[j].tobytes() for j in range(ncol)}) == ncol             # every column has its own randomness
    # domain separation of the nonce: the zero column never reuses a matrix column's stream, whatever stream_id the caller picks
    z = _samples(be, SEED, 0, zero_column=True)
    assert not np.array_equal(z, D[0])
    zero_ct = be.EncryptZero(SEED, streamId=0).download()[0]
    col0_of_zeros = be.EncryptColumns(2048, 4, SEED, count=1, values=np.zeros((1, 2048), np.uint64)).download()[0]
    assert not np.array_equal(zero_ct, col0_of_zeros)
    # the dump IS what the encryptor used: in the NTT domain of limb 0, c0 + c1 s == u e_pk + e0 + e1 s with e_pk = pk0 + pk1 s
    q = int(P.q[0])
    ntt = lambda a: P.ntt(np.array([int(x) % q for x in a], dtype=np.uint64), 0).astype(object)   # noqa: E731
    s_h = ntt(S.sk)
    e_pk = (S.pk[0, 0].astype(object) + S.pk[1, 0].astype(object) * s_h) % q
    for ident, zc, ct in ((0, True, zero_ct), (0, False, col0_of_zeros)):
        uu, a0, a1 = _samples(be, SEED, ident, zero_column=zc)
        lhs = (ct[0, 0].astype(object) + ct[1, 0].astype(object) * s_h) % q
        rhs = (ntt(uu) * e_pk + ntt(a0) + ntt(a1) * s_h) % q
        assert (lhs == rhs).all(), "dumped samples are not the ones inside the ciphertext (zero_column=%s)" % zc
    be.Close()


def test_encrypted_matrix_runs_the_whole_protocol():
    """cmd/server/main.go:187-262 with the matrix encrypted on the device: Encrypt -> Commit -> Prove -> client Decrypt (device) ->
    the checks of Proof.Verify, against the plaintext prover"""
    from lumenos_b200 import fhe
    logN, rows, cols, k = 8, 256, 16, 6
    S = Scenario(logN, rows, cols, k)
    P = S.P
    be = _server(S)
    M = O.witness_matrix(rows, cols)
    matrix = be.EncryptColumns(rows, cols, SEED)
    zero = be.EncryptZero(SEED)
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    lf = fhe.LeafFormat.placeholder(1, P.N)
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    idx = (np.random.default_rng(5).integers(0, 2**62, size=committer.Queries) % (2 * cols)).astype(np.uint32)
    rpt, bpt = be.Encode(S.r), be.Encode(S.b)                      # Prove's plaintexts encoded on the device
    assert np.array_equal(rpt, S.rpt) and np.array_equal(bpt, S.bpt)
    proof = prover.Prove(S.r, S.b, idx, be)                        # slot values: encoded on the device inside Prove
    params = fhe.BGVParameters(logN, P.q, P.p, P.t, P.psi_q, P.psi_p, P.psi_t)
    client = fhe.ClientBFV(fhe.PrimeField(P.t, S.roots), params, sk_as_lattigo(P, S.sk))
    dec = proof.Decrypt(client)
    Mo = M.astype(object)
    r = np.array([int(x) % T for x in S.r], dtype=object)
    b = np.array([int(x) for x in S.b], dtype=object)
    assert [int(x) for x in dec.MatR] == [int(x) % T for x in r.dot(Mo)]
    assert [int(x) for x in dec.MatZ] == [int(x) % T for x in b.dot(Mo)]
    encR, encZ = O.core_encode(dec.MatR, 2, S.roots), O.core_encode(dec.MatZ, 2, S.roots)
    for i, ix in enumerate(idx):
        col = dec.QueriedCols[i].astype(object)
        assert int(col.dot(r)) % T == int(encR[ix]) and int(col.dot(b)) % T == int(encZ[ix]), "well-formedness check failed for column %d" % ix
    client.Close(); be.Close()


@pytest.mark.parametrize("logN,rows,cols,k", [(6, 32, 16, 6), (8, 256, 8, 5), (12, 4096, 8, 4), (14, 16384, 4, 3)])
def test_decrypt_decode_matches_oracle(logN, rows, cols, k):
    from lumenos_b200 import fhe
    S = Scenario(logN, rows, cols, k)
    P = S.P
    params = fhe.BGVParameters(logN, P.q, P.p, P.t, P.psi_q, P.psi_p, P.psi_t)
    client = fhe.ClientBFV(fhe.PrimeField(P.t, S.roots), params, sk_as_lattigo(P, S.sk))
    for g, e in zip(S.gal, S.evks):
        client.LoadGaloisKey(g, e)
    # max-level inputs: every slot
    st = client.NewStore(cols).upload(S.cts)
    got = client.DecryptDecode(st, P.N)
    for j in range(cols):
        assert np.array_equal(got[j], P.decrypt(S.sk, S.cts[j])), j
    assert np.array_equal(got[:, :rows], O.witness_matrix(rows, cols).T)
    # level-1 outputs of the prover with the tracked Scale (Rescale multiplies it by q_l^-1 mod t)
    matR = fhe.matrixInnerSumEval(st, S.rpt, rows, client)
    lvl1 = matR.download()
    got = client.DecryptDecode(matR, 3)
    for j in range(cols):
        assert np.array_equal(got[j], P.decrypt(S.sk, lvl1[j], matR.scale)[:3]), j
    # ranges and the host-array path
    assert np.array_equal(client.DecryptDecode(matR, 1, first=1, count=2), got[1:3, :1])
    assert np.array_equal(client.DecryptArrays(lvl1, matR.scale, 3), got)
    client.Close()


@pytest.mark.parametrize("rows,cols,logN,k,p1", [(8192, 4096, 13, 12, 125815544481056462), (16384, 4096, 14, 12, 5538402014578059)])
def test_config_C_D_whole_protocol_on_real_encryptions(rows, cols, logN, k, p1):
    """BASELINE.json configs[2] and configs[3] (the headline shape) end to end on REAL ciphertexts, which the device encryptor makes
    affordable: the ChaCha20 witness of the reference (P(1) known answer from results/baseline/client/bench_*.txt:22) is encrypted on
    the GPU under an oracle-generated key pair, Commit (Encode + leaves + Merkle tree) and Prove (both inner products with the
    row-swap step, 309 queried columns) run at full size with real Galois keys, the client decrypts on the GPU, and the checks of the
    reference's own test pass: MatR / MatZ == LigeroProveReference (fhe/ligero_test.go:164-174), both well-formedness relations of
    Proof.Verify on every queried column (fhe/ligero.go:554-571), every Merkle path, the evaluation claim.  Two MatR columns and one
    queried column are also compared bit for bit with the oracle evaluated on the downloaded input ciphertexts."""
    import hashlib
    from lumenos_b200 import fhe
    assert O.lib().orc_num_q_primes_for_ntt(cols, T) == k
    P = O.Params(logN, k=k)
    sk = P.keygen_sk(2); pk = P.keygen_pk(sk, 3)
    gal = P.galois_elements_for_innersum(rows)
    evks = [P.keygen_galois(sk, g, 5) for g in gal]
    roots = O.field_roots_forward(2 * cols)
    params = fhe.BGVParameters(logN, P.q, P.p, P.t, P.psi_q, P.psi_p, P.psi_t)
    be = fhe.ClientBFV(fhe.PrimeField(P.t, roots), params, sk_as_lattigo(P, sk))      # server and client roles on one context
    be.LoadPublicKey(pk_as_lattigo(P, pk))
    for g, e in zip(gal, evks):
        be.LoadGaloisKey(g, e)
    be.SetBatch(592)
    M = O.witness_matrix(rows, cols)
    total = int((M & np.uint64(0xFFFFFFFF)).sum(dtype=np.uint64)) + (int((M >> np.uint64(32)).sum(dtype=np.uint64)) << 32)
    assert total % T == p1
    matrix = be.EncryptColumns(rows, cols, SEED)
    zero = be.EncryptZero(SEED)
    committer = fhe.LigeroCommitter(128, rows, cols, 2)
    assert committer.Queries == 309
    lf = fhe.LeafFormat.placeholder(1, P.N)
    prover = committer.Commit(matrix, be, (zero, 0), leafFormat=lf)
    rng = np.random.default_rng(1)
    r = rng.integers(0, 2**63, size=rows, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    z = 3
    zc = pow(z, cols, T)
    b = np.array([pow(zc, i, T) for i in range(rows)], dtype=np.uint64)
    idx = (np.random.default_rng(5).integers(0, 2**62, size=309) % (2 * cols)).astype(np.uint32)
    idx[7] = idx[3]
    proof = prover.Prove(be.Encode(r), be.Encode(b), idx, be)
    dec = proof.Decrypt(be)
    wantR, wantZ = O.vec_mat_mod(r, M), O.vec_mat_mod(b, M)
    assert np.array_equal(dec.MatR, wantR), "MatR differs from the plaintext prover"
    assert np.array_equal(dec.MatZ, wantZ), "MatZ differs from the plaintext prover"
    encR, encZ = O.core_encode(dec.MatR, 2, roots), O.core_encode(dec.MatZ, 2, roots)
    qcols = proof.QueriedCols.download()
    for i, ix in enumerate(idx):
        col = dec.QueriedCols[i]
        assert int(O.vec_mat_mod(r, col[:, None])[0]) == int(encR[ix]), "well-formedness R check failed for column %d" % ix
        assert int(O.vec_mat_mod(b, col[:, None])[0]) == int(encZ[ix]), "well-formedness B check failed for column %d" % ix
        leaf = O.leaf_bytes(qcols[i], lf.header, lf.polyHeader, lf.limbHeader)
        assert fhe.VerifyMerklePath(leaf, proof.MerklePaths[i], proof.Root, int(ix)), "merkle path of column %d" % ix
    a = np.array([pow(z, i, T) for i in range(cols)], dtype=np.uint64)
    assert int(O.vec_mat_mod(a, dec.MatZ[:, None])[0]) == int(O.vec_mat_mod(a, wantZ[:, None])[0])      # the evaluation claim
    # P1 on samples: the oracle evaluated on the very ciphertexts the device encryptor produced
    rpt = be.Encode(r)
    for j in (0, cols - 1):
        want, _ = P.inner_sum_eval_column(matrix.download(j, 1)[0], rpt, rows, gal, evks)
        assert np.array_equal(proof.MatR.download(j, 1)[0], want), "MatR[%d] differs from the oracle" % j
    assert np.array_equal(qcols[3], qcols[7])
    assert np.array_equal(qcols[0], P.rescale_to_level1(prover.EncodedMatrix.download(int(idx[0]), 1)[0])[0])
    be.Close()


def test_client_entry_points_fail_loudly():
    """error behaviour of the new entry points: missing ring / keys / communicator and bad shapes return a status, never abort"""
    import ctypes as C
    from lumenos_b200 import _capi, fhe
    S = Scenario(6, 32, 8, 3)
    be = S.backend()
    lib = _capi.lib()
    st = be.NewStore(8)
    with pytest.raises(fhe.LumenosError, match="plaintext ring not loaded"):
        check_rc = lib.lmn_encrypt_columns(be.ctx, None, fhe.REFERENCE_WITNESS_KEY, 32, 8, 0, 8, SEED, st.h)
        _capi.check(check_rc)
    be.LoadPlaintextRing()
    with pytest.raises(fhe.LumenosError, match="public key not loaded"):
        be.EncryptColumns(32, 8, SEED, out=st)
    with pytest.raises(fhe.LumenosError, match="secret key not loaded"):
        out = np.empty((8, 4), np.uint64)
        _capi.check(lib.lmn_decrypt_decode(be.ctx, st.h, 0, 8, 1, 4, out.ctypes.data_as(C.c_void_p)))
    be.LoadPublicKey(pk_as_lattigo(S.P, S.pk))
    with pytest.raises(fhe.LumenosError, match="column range"):
        be.EncryptColumns(32, 8, SEED, firstCol=4, count=8, out=st)
    with pytest.raises(fhe.LumenosError, match="between 1 and N"):
        be.EncryptColumns(65, 8, SEED, out=st)
    with pytest.raises(fhe.LumenosError, match="max-level"):
        be.EncryptColumns(32, 8, SEED, out=be.NewStore(8, 1))
    with pytest.raises(fhe.LumenosError, match="primitive 2N-th root"):
        be.LoadPlaintextRing(12345)
    with pytest.raises(fhe.LumenosError, match="no communicator"):
        _capi.check(lib.lmn_exchange_encoded(be.ctx, C.c_void_p(st.device_ptr), 8, st.h))
    with pytest.raises(fhe.LumenosError, match="power of two"):
        _capi.check(lib.lmn_comm_init(be.ctx, bytes(128), 0, 3))
    with pytest.raises(fhe.LumenosError, match="Scale must be invertible"):
        out = np.empty((8, 4), np.uint64)
        _capi.check(lib.lmn_decrypt_decode(be.ctx, st.h, 0, 8, 0, 4, out.ctypes.data_as(C.c_void_p)))
    # a single-rank communicator works and the exchanges degenerate to copies
    from lumenos_b200 import sharding
    import torch
    sharding.init_comm(be, 0, 1)
    be.EncryptColumns(32, 8, SEED, out=st)
    t = torch.empty((8, 2, 3, 64), dtype=torch.int64, device="cuda")
    sharding.to_slices(be, st, 8, t)
    be.Synchronize()
    assert np.array_equal(t.cpu().numpy().astype(np.uint64), st.download())
    blk = sharding.exchange_encoded(be, t)
    assert np.array_equal(blk.download(), st.download())
    # the exchange entry points refuse a store of another context (as the client entry points do)
    be2 = S.backend()
    foreign = be2.NewStore(8)
    tp = C.c_void_p(t.data_ptr())
    idx = (C.c_uint32 * 2)(0, 1)
    for rc in (lib.lmn_exchange_encoded(be.ctx, tp, 8, foreign.h), lib.lmn_exchange_to_slices(be.ctx, foreign.h, 8, tp),
               lib.lmn_exchange_queried(be.ctx, tp, 8, idx, 2, foreign.h), lib.lmn_commit_leaves_from_slices(be.ctx, tp, 8, be2.NewStore(8, 1).h)):
        assert rc == -1 and "different contexts" in lib.lmn_last_error().decode()
    be2.Close()
    be.Close()
