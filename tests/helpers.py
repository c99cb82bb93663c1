"""Shared scenario builder for the parity tests.  The oracle plays the role of the Go/Lattigo host: it
generates parameters, keys, plaintexts and ciphertexts, which are handed to the product through the C ABI
exactly as the cgo shim would."""
import numpy as np

from oracle import oracle as O


class Scenario:
    def __init__(self, logN, rows, cols, k, rho_inv=2, seed=1, switches=None):
        self.O = O
        self.P = O.Params(logN, k=k, switches=switches)
        self.logN, self.rows, self.cols, self.rho_inv = logN, rows, cols, rho_inv
        P = self.P
        self.sk = P.keygen_sk(seed + 1)
        self.pk = P.keygen_pk(self.sk, seed + 2)
        self.M = O.witness_matrix(rows, cols)                       # row-major rows x cols
        self.roots = O.field_roots_forward(cols * rho_inv)
        self.cts = np.stack([P.encrypt(self.pk, P.encode(self.M[:, j]), 1000 + j) for j in range(cols)])
        self.zero = P.encrypt(self.pk, None, seed + 3)
        self.gal = P.galois_elements_for_innersum(rows)
        self.evks = [P.keygen_galois(self.sk, g, seed + 4) for g in self.gal]
        rng = np.random.default_rng(seed)
        self.r = rng.integers(0, 2**63, size=rows, dtype=np.uint64) * np.uint64(2) + np.uint64(1)  # unreduced u64, like SampleUints
        z = 3
        zc = pow(z, cols, P.t)
        self.b = np.array([pow(zc, i, P.t) for i in range(rows)], dtype=np.uint64)                     # ligero.go:210-216
        self.z = z
        self.rpt, self.bpt = P.encode(self.r), P.encode(self.b)

    def backend(self, **kw):
        from lumenos_b200 import fhe
        P = self.P
        params = fhe.BGVParameters(self.logN, P.q, P.p, P.t, P.psi_q, P.psi_p, P.psi_t)
        field = fhe.PrimeField(P.t, self.roots)
        be = fhe.ServerBFV(field, params, **kw)
        for g, e in zip(self.gal, self.evks):
            be.LoadGaloisKey(g, e)
        return be


def mform(x, q):
    """x * 2^64 mod q elementwise (Lattigo keeps keys in Montgomery form)"""
    return np.array([(int(v) << 64) % int(q) for v in np.asarray(x).ravel()], dtype=np.uint64).reshape(np.shape(x))


def pk_as_lattigo(P, pk):
    """oracle public key [2][k][N] (NTT, canonical) -> NTT + Montgomery, as rlwe.PublicKey holds it"""
    out = np.empty_like(pk)
    for i in range(P.k):
        out[:, i, :] = mform(pk[:, i, :], P.q[i])
    return out


def sk_as_lattigo(P, sk):
    """oracle secret key (ternary coefficients) -> [k][N] NTT + Montgomery, as rlwe.SecretKey.Value.Q holds it"""
    out = np.empty((P.k, P.N), np.uint64)
    for i in range(P.k):
        q = int(P.q[i])
        out[i] = mform(P.ntt(np.array([int(v) % q for v in sk], dtype=np.uint64), i), q)
    return out
