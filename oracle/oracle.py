"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE -- never imported by lumenos_b200).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
See oracle/lumenos_oracle.h for the parity status ("ciphertext-level parity unpinned").
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblumenos_oracle.so")

T_DEFAULT = 144115188075593729  # 2^57 - 2^18 + 1, ref: cmd/server/main.go:22, fhe/ligero_test.go:16


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("lumenos_oracle.c", "merkle_oracle.c", "lumenos_oracle.h")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


class Switches(C.Structure):
    _fields_ = [("primroot_start", C.c_int), ("scalar_centered", C.c_int), ("digit_centered", C.c_int),
                ("moddown_round_nearest", C.c_int), ("innersum_rowswap_last", C.c_int)]


_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(build())
    vp, i, u64 = C.c_void_p, C.c_int, C.c_uint64
    sig = {
        "orc_default_switches": (None, [C.POINTER(Switches)]),
        "orc_sha256": (None, [C.c_char_p, u64, C.c_char_p]),
        "orc_leaf_bytes": (u64, [_u64p, i, i, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_char_p]),
        "orc_proof_bytes": (u64, [C.c_uint32] * 4 + [_u64p, _u64p, C.c_uint32, i, i] + [C.c_char_p, C.c_uint32] * 3 + [_u64p, C.c_uint32, i, i]
                            + [C.c_char_p, C.c_uint32] * 3 + [C.c_char_p, u64, C.c_char_p, C.c_uint32, C.c_char_p]),
        "orc_merkle_tree": (C.c_uint32, [C.c_char_p, C.c_uint32, C.c_char_p]),
        "orc_merkle_path": (C.c_uint32, [C.c_char_p, C.c_uint32, C.c_uint32, C.c_char_p]),
        "orc_params_generate": (vp, [i, i, u64, C.POINTER(Switches)]),
        "orc_params_from_moduli": (vp, [i, i, _u64p, i, _u64p, u64, C.POINTER(Switches)]),
        "orc_params_free": (None, [vp]),
        "orc_params_k": (i, [vp]), "orc_params_np": (i, [vp]), "orc_params_N": (i, [vp]),
        "orc_params_t": (u64, [vp]),
        "orc_params_moduli": (None, [vp, _u64p, _u64p]),
        "orc_params_psi": (None, [vp, _u64p, _u64p]),
        "orc_params_psi_t": (u64, [vp]),
        "orc_num_q_primes_for_ntt": (i, [i, u64]),
        "orc_witness_matrix": (None, [i, i, u64, _u64p]),
        "orc_field_roots_forward": (i, [u64, i, i, _u64p]),
        "orc_vec_mat_mod": (None, [_u64p, _u64p, i, i, u64, _u64p]),
        "orc_core_ntt": (None, [_u64p, i, i, u64, _u64p, i]),
        "orc_core_encode": (None, [_u64p, i, i, u64, _u64p, _u64p]),
        "orc_ntt_limb": (None, [vp, i, _u64p]), "orc_intt_limb": (None, [vp, i, _u64p]),
        "orc_automorphism_index": (None, [vp, u64, _u32p]),
        "orc_keygen_sk": (None, [vp, u64, _i8p]),
        "orc_keygen_pk": (None, [vp, _i8p, u64, _u64p]),
        "orc_beta": (i, [vp]),
        "orc_keygen_galois": (None, [vp, _i8p, u64, u64, _u64p]),
        "orc_keygen_ringswitch": (None, [vp, _i8p, _i8p, i, u64, _u64p]),
        "orc_encode_pt": (None, [vp, _u64p, i, i, _u64p]),
        "orc_encrypt_pk": (None, [vp, _u64p, C.c_void_p, u64, _u64p]),
        "orc_decrypt_decode": (None, [vp, _i8p, _u64p, i, u64, _u64p]),
        "orc_noise_bits": (C.c_double, [vp, _i8p, _u64p, i]),
        "orc_ct_add": (None, [vp, i, _u64p, _u64p, _u64p]),
        "orc_ct_sub": (None, [vp, i, _u64p, _u64p, _u64p]),
        "orc_ct_mul_scalar": (None, [vp, i, _u64p, u64, _u64p]),
        "orc_ct_mul_pt": (None, [vp, i, _u64p, _u64p, _u64p]),
        "orc_galois_elements_for_innersum": (i, [vp, i, _u64p]),
        "orc_ct_innersum": (i, [vp, i, _u64p, i, _u64p, C.POINTER(C.c_void_p), i]),
        "orc_ct_rescale": (u64, [vp, i, _u64p, _u64p, u64]),
        "orc_ct_automorphism": (None, [vp, i, _u64p, u64, _u64p, _u64p]),
        "orc_ct_ringswitch": (None, [vp, _u64p, _u64p, i, _u64p]),
        "orc_fhe_ntt": (None, [vp, _u64p, i, i, _u64p, i]),
        "orc_fhe_encode": (None, [vp, _u64p, i, i, _u64p, _u64p, _u64p]),
        "orc_fhe_encode_lanes": (None, [vp, _u64p, i, i, _u64p, _i32p, i, _u64p, _u64p]),
        "orc_inner_sum_eval_column": (u64, [vp, _u64p, _u64p, i, _u64p, C.POINTER(C.c_void_p), i, _u64p]),
        "orc_rescale_to_level1": (u64, [vp, _u64p, _u64p]),
        "orc_fhe_ntt_counts": (None, [C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    _lib = L
    return L


def default_switches(**kw):
    sw = Switches()
    lib().orc_default_switches(C.byref(sw))
    for k, v in kw.items():
        setattr(sw, k, v)
    return sw


class Params:
    """BGV parameters as produced by fhe.GenerateBGVParamsForNTT (ref: fhe/bfv.go:121-188)."""

    def __init__(self, logN, k=None, t=T_DEFAULT, cols=None, q=None, p=None, switches=None):
        L = lib()
        self.sw = switches if switches is not None else default_switches()
        if q is not None:
            q = np.ascontiguousarray(q, dtype=np.uint64)
            p = np.ascontiguousarray(p, dtype=np.uint64)
            self.h = L.orc_params_from_moduli(logN, len(q), q, len(p), p, t, C.byref(self.sw))
        else:
            if k is None:
                k = L.orc_num_q_primes_for_ntt(cols, t)
            self.h = L.orc_params_generate(logN, k, t, C.byref(self.sw))
        if not self.h:
            raise ValueError("invalid BGV parameters (t must be 1 mod 2N)")
        self.logN, self.N, self.t = logN, 1 << logN, t
        self.k, self.np = L.orc_params_k(self.h), L.orc_params_np(self.h)
        self.q = np.zeros(self.k, np.uint64)
        self.p = np.zeros(self.np, np.uint64)
        L.orc_params_moduli(self.h, self.q, self.p)
        self.psi_q = np.zeros(self.k, np.uint64)
        self.psi_p = np.zeros(self.np, np.uint64)
        L.orc_params_psi(self.h, self.psi_q, self.psi_p)
        self.psi_t = L.orc_params_psi_t(self.h)
        self.beta = L.orc_beta(self.h)
        self.max_level = self.k - 1

    def __del__(self):
        try:
            if self.h:
                lib().orc_params_free(self.h)
                self.h = None
        except Exception:
            pass

    # -- sizes --
    def ct_words(self, level=None):
        level = self.max_level if level is None else level
        return 2 * (level + 1) * self.N

    def new_ct(self, level=None, count=None):
        level = self.max_level if level is None else level
        shape = (2, level + 1, self.N) if count is None else (count, 2, level + 1, self.N)
        return np.zeros(shape, np.uint64)

    # -- client side --
    def keygen_sk(self, seed):
        sk = np.zeros(self.N, np.int8)
        lib().orc_keygen_sk(self.h, seed, sk)
        return sk

    def keygen_pk(self, sk, seed):
        pk = np.zeros((2, self.k, self.N), np.uint64)
        lib().orc_keygen_pk(self.h, sk, seed, pk)
        return pk

    def keygen_galois(self, sk, galEl, seed):
        evk = np.zeros((self.beta, 2, self.k + self.np, self.N), np.uint64)
        lib().orc_keygen_galois(self.h, sk, int(galEl), seed, evk)
        return evk

    def keygen_ringswitch(self, sk, sk_small, logNs, seed):
        evk = np.zeros((self.beta, 2, self.k + self.np, self.N), np.uint64)
        lib().orc_keygen_ringswitch(self.h, sk, sk_small, logNs, seed, evk)
        return evk

    def galois_elements_for_innersum(self, n):
        g = np.zeros(64, np.uint64)
        c = lib().orc_galois_elements_for_innersum(self.h, n, g)
        return [int(x) for x in g[:c]]

    def encode(self, values, level=None):
        level = self.max_level if level is None else level
        values = np.ascontiguousarray(values, dtype=np.uint64)
        pt = np.zeros((level + 1, self.N), np.uint64)
        lib().orc_encode_pt(self.h, values, len(values), level, pt)
        return pt

    def encrypt(self, pk, pt, seed):
        ct = self.new_ct()
        ptr = None if pt is None else pt.ctypes.data_as(C.c_void_p)
        lib().orc_encrypt_pk(self.h, pk, ptr, seed, ct)
        return ct

    def decrypt(self, sk, ct, scale=1):
        level = ct.shape[-2] - 1
        out = np.zeros(self.N, np.uint64)
        lib().orc_decrypt_decode(self.h, sk, np.ascontiguousarray(ct), level, int(scale), out)
        return out

    def noise_bits(self, sk, ct):
        return lib().orc_noise_bits(self.h, sk, np.ascontiguousarray(ct), ct.shape[-2] - 1)

    # -- ring --
    def ntt(self, a, mod_idx):
        a = np.ascontiguousarray(a, np.uint64).copy()
        lib().orc_ntt_limb(self.h, mod_idx, a)
        return a

    def intt(self, a, mod_idx):
        a = np.ascontiguousarray(a, np.uint64).copy()
        lib().orc_intt_limb(self.h, mod_idx, a)
        return a

    def automorphism_index(self, galEl):
        idx = np.zeros(self.N, np.uint32)
        lib().orc_automorphism_index(self.h, int(galEl), idx)
        return idx

    # -- ServerBFV operator level --
    def add(self, a, b):
        out = np.empty_like(a); lib().orc_ct_add(self.h, a.shape[-2] - 1, a, b, out); return out

    def sub(self, a, b):
        out = np.empty_like(a); lib().orc_ct_sub(self.h, a.shape[-2] - 1, a, b, out); return out

    def mul_scalar(self, a, w):
        out = np.empty_like(a); lib().orc_ct_mul_scalar(self.h, a.shape[-2] - 1, a, int(w), out); return out

    def mul_pt(self, a, pt):
        out = np.empty_like(a); lib().orc_ct_mul_pt(self.h, a.shape[-2] - 1, a, np.ascontiguousarray(pt), out); return out

    @staticmethod
    def _keyptrs(evks):
        arr = (C.c_void_p * len(evks))(*[e.ctypes.data for e in evks])
        return arr

    def innersum(self, ct, n, galEls, evks):
        ct = np.ascontiguousarray(ct).copy()
        g = np.ascontiguousarray(galEls, dtype=np.uint64)
        rc = lib().orc_ct_innersum(self.h, ct.shape[-2] - 1, ct, n, g, self._keyptrs(evks), len(evks))
        if rc != 0:
            raise RuntimeError("oracle InnerSum failed: %d" % rc)
        return ct

    def rescale(self, ct, scale=1):
        level = ct.shape[-2] - 1
        out = np.zeros((2, level, self.N), np.uint64)
        s = lib().orc_ct_rescale(self.h, level, np.ascontiguousarray(ct), out, int(scale))
        return out, int(s)

    def rescale_to_level1(self, ct):
        out = np.zeros((2, min(2, self.k), self.N), np.uint64)
        s = lib().orc_rescale_to_level1(self.h, np.ascontiguousarray(ct), out)
        return out, int(s)

    def automorphism(self, ct, galEl, evk):
        out = np.empty_like(ct)
        lib().orc_ct_automorphism(self.h, ct.shape[-2] - 1, np.ascontiguousarray(ct), int(galEl), evk, out)
        return out

    def ringswitch(self, ct_lvl1, evk, logNs):
        out = np.zeros((2, 1, 1 << logNs), np.uint64)
        lib().orc_ct_ringswitch(self.h, np.ascontiguousarray(ct_lvl1), evk, logNs, out)
        return out

    # -- path level --
    def fhe_ntt(self, cts, size, roots):
        cts = np.ascontiguousarray(cts).copy()
        lib().orc_fhe_ntt(self.h, cts, cts.shape[0], size, roots, len(roots))
        return cts

    def fhe_encode(self, matrix, rho_inv, zero_ct, roots):
        cols = matrix.shape[0]
        enc = np.zeros((cols * rho_inv,) + matrix.shape[1:], np.uint64)
        lib().orc_fhe_encode(self.h, np.ascontiguousarray(matrix), cols, rho_inv, np.ascontiguousarray(zero_ct), roots, enc)
        return enc

    def fhe_encode_lanes(self, matrix_lanes, rho_inv, zero_lanes, limb_of_lane, roots):
        cols, nl = matrix_lanes.shape
        enc = np.zeros((cols * rho_inv, nl), np.uint64)
        lib().orc_fhe_encode_lanes(self.h, np.ascontiguousarray(matrix_lanes), cols, rho_inv,
                                   np.ascontiguousarray(zero_lanes), np.ascontiguousarray(limb_of_lane, dtype=np.int32),
                                   nl, roots, enc)
        return enc

    def inner_sum_eval_column(self, ct, pt, rows, galEls, evks):
        out = np.zeros((2, 2, self.N), np.uint64)
        g = np.ascontiguousarray(galEls, dtype=np.uint64)
        s = lib().orc_inner_sum_eval_column(self.h, np.ascontiguousarray(ct), np.ascontiguousarray(pt), rows, g,
                                            self._keyptrs(evks), len(evks), out)
        if s == 0:
            raise RuntimeError("oracle matrixInnerSumEval failed (missing Galois key or unsupported rows)")
        return out, int(s)


def ntt_counts():
    a, s, m = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().orc_fhe_ntt_counts(C.byref(a), C.byref(s), C.byref(m))
    return a.value, s.value, m.value


def witness_matrix(rows, cols, t=T_DEFAULT):
    """core.RandomMatrixRowMajor (ref: core/utils.go:46-82): row-major rows x cols."""
    out = np.zeros((rows, cols), np.uint64)
    lib().orc_witness_matrix(rows, cols, t, out)
    return out


def field_roots_forward(n_field, t=T_DEFAULT, primroot_start=3):
    """RootsForward table of core.NewPrimeField(t, n_field) (ref: core/field.go:138-197)."""
    tab = np.zeros(n_field, np.uint64)
    if lib().orc_field_roots_forward(t, n_field, primroot_start, tab) != 0:
        raise ValueError("invalid field parameters")
    return tab


def vec_mat_mod(r, M, t=T_DEFAULT):
    """sum_i r_i M[i][j] mod t for every column j of the row-major matrix M (the plaintext prover's MatR / MatZ)"""
    M = np.ascontiguousarray(M, np.uint64)
    out = np.zeros(M.shape[1], np.uint64)
    lib().orc_vec_mat_mod(np.ascontiguousarray(r, np.uint64), M, M.shape[0], M.shape[1], t, out)
    return out


def core_ntt(values, size, roots, t=T_DEFAULT):
    v = np.ascontiguousarray(values, np.uint64).copy()
    lib().orc_core_ntt(v, len(v), size, t, roots, len(roots))
    return v


def core_encode(row, rho_inv, roots, t=T_DEFAULT):
    row = np.ascontiguousarray(row, np.uint64)
    out = np.zeros(len(row) * rho_inv, np.uint64)
    lib().orc_core_encode(row, len(row), rho_inv, t, roots, out)
    return out


# ---- "Merkle tree built" span of Commit (merkle_oracle.c) ----
def sha256(data):
    out = C.create_string_buffer(32)
    lib().orc_sha256(bytes(data), len(data), out)
    return out.raw


def leaf_bytes(ct, header=b"", poly_header=b"", limb_header=b""):
    """ct: [2][nl][N] level-1 ciphertext -> the leaf buffer of processLeafParallel (fhe/ligero.go:157-158)"""
    ct = np.ascontiguousarray(ct, dtype=np.uint64)
    nl, N = ct.shape[1], ct.shape[2]
    n = len(header) + 2 * (len(poly_header) + nl * (len(limb_header) + 8 * N))
    out = C.create_string_buffer(n)
    got = lib().orc_leaf_bytes(ct, nl, N, header, len(header), poly_header, len(poly_header), limb_header, len(limb_header), out)
    assert got == n
    return out.raw


def proof_bytes(rows, cols, rho_inv, queries, mat_r, mat_z, queried, mat_fmt, query_fmt=None, paths=(), root=b""):
    """EncryptedProof.WriteTo (fhe/ligero.go:660-704).  mat_r / mat_z: [cols][2][nl][N]; queried: [nq][2][nl'][N']; *_fmt = (header,
    poly_header, limb_header); paths: list of lists of 32-byte hashes."""
    mat_r, mat_z = np.ascontiguousarray(mat_r, dtype=np.uint64), np.ascontiguousarray(mat_z, dtype=np.uint64)
    queried = np.ascontiguousarray(queried, dtype=np.uint64)
    qf = query_fmt or mat_fmt
    pb = b"".join(b"".join(p) for p in paths)
    fm = [x for h in mat_fmt for x in (bytes(h), len(h))]
    fq = [x for h in qf for x in (bytes(h), len(h))]
    args = [rows & 0xFFFFFFFF, cols & 0xFFFFFFFF, rho_inv & 0xFFFFFFFF, queries & 0xFFFFFFFF, mat_r, mat_z, mat_r.shape[0], mat_r.shape[2],
            mat_r.shape[3]] + fm + [queried, queried.shape[0], queried.shape[2], queried.shape[3]] + fq + [pb, len(pb), bytes(root), len(root)]
    n = lib().orc_proof_bytes(*args, None)
    out = C.create_string_buffer(max(1, n))
    assert lib().orc_proof_bytes(*args, out) == n
    return out.raw[:n]


class MerkleTree:
    """core.NewTree over leaf buffers (tree.go:76-163)"""

    def __init__(self, leaf_buffers):
        self.n = len(leaf_buffers)
        hashes = b"".join(sha256(b) for b in leaf_buffers)
        total = lib().orc_merkle_tree(hashes, self.n, None)
        self.nodes = C.create_string_buffer(32 * max(1, total))
        lib().orc_merkle_tree(hashes, self.n, self.nodes)
        self.total = total

    def root(self):
        return self.nodes.raw[32 * (self.total - 1):32 * self.total]

    def leaf_hash(self, i):
        return self.nodes.raw[32 * i:32 * i + 32]

    def path(self, index):
        out = C.create_string_buffer(32 * 40)
        d = lib().orc_merkle_path(self.nodes, self.n, index, out)
        return [out.raw[32 * j:32 * j + 32] for j in range(d)]
