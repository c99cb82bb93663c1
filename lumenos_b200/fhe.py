"""Host-side mirror of the reference's `fhe` package surface for the accelerated path, over the C ABI.

Same names, argument meaning and error behaviour as the Go interface (ChainSafe/lumenos @ 9f33d62):
    GenerateBGVParamsForNTT   fhe/bfv.go:121-188        ServerBFV            fhe/bfv.go:13-58
    Encode                    fhe/code.go:8-34          NTT                  fhe/ntt.go:12-17
    LigeroCommitter.Commit    fhe/ligero.go:95-124      LigeroProver.Prove   fhe/ligero.go:194-291
    matrixInnerSumEval        fhe/ligero.go:299-370     RingSwitchServer     fhe/ring_switch.go:93-113
What stays on the Go host in the reference (keys, encryption, the merlin transcript, Merkle tree, wire
format) is passed in as data: plaintexts, Galois keys, the zero-column ciphertext and query indices.
Ciphertexts live in device-resident stores; numpy is used only to carry host buffers into ctypes.
"""
import ctypes as C
import math
import weakref

import numpy as np

from . import _capi
from ._capi import LumenosError, check  # noqa: F401

DEFAULT = 0xFFFFFFFF


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _u64(a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return a


def GenerateBGVParamsForNTT(nttSize, logN, plaintextModulus):
    """ref: fhe/bfv.go:121-188.  Returns the bgv.ParametersLiteral fields (prime generation is the host's)."""
    if nttSize < 2:
        raise ValueError("nttSize must be >= 2")
    if logN <= 0:
        raise ValueError("logN must be positive")
    two_n = 2 << logN
    if plaintextModulus % two_n != 1:
        raise ValueError("plaintextModulus T (%d) does not satisfy T = 1 (mod 2N) (2N=%d)" % (plaintextModulus, two_n))
    buffer_levels = 0 if plaintextModulus.bit_length() > 45 else -2
    k = (nttSize & -nttSize).bit_length() - 1 + buffer_levels
    return {"LogN": logN, "LogQ": [58] + [56] * (k - 1), "LogP": [55, 55], "PlaintextModulus": plaintextModulus}


def calculateQueries(securityBits, rhoInv):
    """ref: fhe/ligero.go:65-71"""
    term = math.log2(1.0 + 1.0 / rhoInv)
    if 1.0 - term <= 0:
        return 0
    return int(math.ceil(securityBits / (1.0 - term)))


class BGVParameters:
    """bgv.Parameters as the host holds them: moduli and the primitive 2N-th roots its rings use."""

    def __init__(self, logN, Q, P, T, psi_q, psi_p, psi_t=None):
        self.logN, self.N, self.T = int(logN), 1 << int(logN), int(T)
        self.Q, self.P = _u64(Q), _u64(P)
        self.psi_q, self.psi_p = _u64(psi_q), _u64(psi_p)
        self.psi_t = None if psi_t is None else int(psi_t)   # primitive 2N-th root of the encoder's ring T (Encode / Decode / Encrypt only)

    def MaxLevel(self):
        return len(self.Q) - 1

    def GaloisElement(self, k):
        return pow(5, k % (self.N // 2), 2 * self.N)

    def GaloisElementsForInnerSum(self, batch, n):
        """lattigo bgv.Parameters.GaloisElementsForInnerSum(batch, n) (ref: fhe/ligero_test.go:53, cmd/client/main.go:81) for power-of-two n:
        the rotations batch * {2^i : 2^i < n} and batch * n, plus the row swap when n * batch > N/2.  Rotations by multiples of N/2 are
        the identity (Galois element 1): Lattigo generates keys for them all the same - the logged key-blob sizes count them
        (tests/test_host_logic.py::test_key_blob_sizes_pin_the_key_shapes_and_the_galois_set) - and the library ignores them."""
        g, s = [], 1
        while s < n:
            g.append(self.GaloisElement(s * batch))
            s <<= 1
        g.append(self.GaloisElement(n * batch))
        if n * batch > self.N // 2:
            g.append(2 * self.N - 1)
        return g


class CiphertextStore:
    """Device-resident []*rlwe.Ciphertext (ref: LigeroProver{Matrix, EncodedMatrix}, fhe/ligero.go:32-37)."""

    def __init__(self, backend, count, level=None, small=False, wrap_ptr=None):
        self.backend = backend
        h = C.c_void_p()
        L = _capi.lib()
        if small:
            check(L.lmn_store_create_small(backend.ctx, count, C.byref(h)))
            self.N = 1 << backend.ring_switch_logn
        else:
            level = backend.params.MaxLevel() if level is None else level
            if wrap_ptr is not None:
                check(L.lmn_store_wrap(backend.ctx, wrap_ptr, count, level, C.byref(h)))
            else:
                check(L.lmn_store_create(backend.ctx, count, level, C.byref(h)))
            self.N = backend.params.N
        self.h = h
        backend._stores.add(self)
        self.count = count
        self.level = L.lmn_store_level(h)
        self.scale = 1   # rlwe.MetaData.Scale (mod T), tracked on the host as in the reference

    def __len__(self):
        return self.count

    def shape(self, count=None):
        return (self.count if count is None else count, 2, self.level + 1, self.N)

    def upload(self, cts, first=0):
        cts = _u64(cts)
        n = cts.size // (2 * (self.level + 1) * self.N)
        check(_capi.lib().lmn_store_upload(self.h, first, n, _ptr(cts)))
        return self

    def upload_limb(self, ct, poly, limb, data):
        data = _u64(data)
        check(_capi.lib().lmn_store_upload_limb(self.h, ct, poly, limb, _ptr(data)))

    @staticmethod
    def limb_table(limbs):
        """C pointer table over a sequence of per-limb uint64 arrays in store order (ct, poly, limb): what a Go host builds from
        ct.Value[p].Coeffs[l] (core/utils.go:12-33) after pinning the slices"""
        tab = (C.c_void_p * len(limbs))()
        for i, a in enumerate(limbs):
            if a.dtype != np.uint64 or not a.flags.c_contiguous:
                raise ValueError("every limb must be a contiguous uint64 array")
            tab[i] = a.ctypes.data
        return tab

    def upload_limbs(self, limbs, first=0, count=None, table=None):
        """limbs: count * 2 * (level+1) separate arrays of N uint64 (any host memory) -> ciphertexts [first, first + count)"""
        count = self.count - first if count is None else count
        table = self.limb_table(limbs) if table is None else table
        if len(table) != count * 2 * (self.level + 1):
            raise ValueError("the table needs count * 2 * (level + 1) limbs")
        check(_capi.lib().lmn_store_upload_limbs(self.h, first, count, table))
        return self

    def download_limbs(self, limbs, first=0, count=None, table=None):
        count = self.count - first if count is None else count
        table = self.limb_table(limbs) if table is None else table
        if len(table) != count * 2 * (self.level + 1):
            raise ValueError("the table needs count * 2 * (level + 1) limbs")
        check(_capi.lib().lmn_store_download_limbs(self.h, first, count, table))
        return limbs

    def download(self, first=0, count=None):
        count = self.count - first if count is None else count
        out = np.empty(self.shape(count), np.uint64)
        check(_capi.lib().lmn_store_download(self.h, first, count, _ptr(out)))
        return out

    def download_limb(self, ct, poly, limb):
        out = np.empty(self.N, np.uint64)
        check(_capi.lib().lmn_store_download_limb(self.h, ct, poly, limb, _ptr(out)))
        return out

    @property
    def device_ptr(self):
        return _capi.lib().lmn_store_device_ptr(self.h)

    @property
    def nbytes(self):
        return _capi.lib().lmn_store_bytes(self.h)

    def free(self):
        # a store must be released before its context (ServerBFV.Close releases the survivors itself)
        if self.h and self.backend.ctx:
            _capi.lib().lmn_store_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PrimeField:
    """core.PrimeField as far as the path reads it (ref: core/field.go:29-47): N() and RootForwardUint64(i).
    The table itself is the host's (it comes from Lattigo's ring.SubRing)."""

    def __init__(self, modulus, roots_forward):
        self.modulus = int(modulus)
        self.roots = _u64(roots_forward)

    def N(self):
        return len(self.roots)

    def Modulus(self):
        return self.modulus

    def RootForwardUint64(self, i):
        return int(self.roots[i])


class RingSwitchServer:
    """ref: fhe/ring_switch.go:93-113"""

    def __init__(self, backend, ringSwitchEvk, logN_small, psi_small=0):
        """ringSwitchEvk: evk.Value as [rns digits][2][k+np][N] (one base-2 digit) or [rns digits][base-2 digits][2][k+np][N]"""
        evk = _u64(ringSwitchEvk)
        rns, b2 = (evk.shape[0], 1) if evk.ndim == 4 else (evk.shape[0], evk.shape[1])
        check(_capi.lib().lmn_load_ring_switch_key(backend.ctx, logN_small, _ptr(evk), rns, b2, psi_small))
        backend.ring_switch_logn = logN_small
        self.backend = backend

    def RingSwitchNew(self, store):
        out = CiphertextStore(self.backend, store.count, small=True)
        check(_capi.lib().lmn_ring_switch(self.backend.ctx, store.h, out.h))
        out.scale = store.scale
        return out


class ServerBFV:
    """ref: fhe/bfv.go:13-58 (operator boundary of the path).  One instance = one GPU."""

    def __init__(self, ptField, params, device=0, scalar_centered=DEFAULT, moddown_round_nearest=DEFAULT,
                 innersum_rowswap_last=DEFAULT, stream=None):
        self.ptField, self.params = ptField, params
        p = _capi.LmnParams()
        p.log_n, p.num_q, p.num_p, p.t = params.logN, len(params.Q), len(params.P), params.T
        p.q, p.p = params.Q.ctypes.data_as(_capi.u64p), params.P.ctypes.data_as(_capi.u64p)
        p.psi_q, p.psi_p = params.psi_q.ctypes.data_as(_capi.u64p), params.psi_p.ctypes.data_as(_capi.u64p)
        p.scalar_centered, p.moddown_round_nearest = scalar_centered, moddown_round_nearest
        p.innersum_rowswap_last = innersum_rowswap_last
        h = C.c_void_p()
        check(_capi.lib().lmn_ctx_create(C.byref(p), device, C.byref(h)))
        self.ctx = h
        self.device = device
        self.rs = None
        self.ring_switch_logn = 0
        self.mulCounter = 0
        self._plans = {}
        self._stores = weakref.WeakSet()
        if stream is not None:
            self.SetStream(stream)

    # -- plumbing --
    def SetStream(self, cuda_stream_ptr):
        # torch's default stream has handle 0, which the C ABI reads as "the context's own stream": name the legacy default
        # stream explicitly (cudaStreamLegacy = 0x1) so that events recorded on torch's stream order with the library's kernels
        if cuda_stream_ptr == 0:
            cuda_stream_ptr = 1
        check(_capi.lib().lmn_ctx_set_stream(self.ctx, cuda_stream_ptr))

    def Synchronize(self):
        check(_capi.lib().lmn_ctx_synchronize(self.ctx))

    def SetBatch(self, jobs):
        check(_capi.lib().lmn_ctx_set_batch(self.ctx, jobs))

    def Trim(self):
        """hand the pooled scratch of the context (batch workspaces, serialised leaves, exchange staging) back to the driver;
        returns the bytes released.  Everything regrows on demand."""
        n = C.c_size_t(0)
        check(_capi.lib().lmn_ctx_trim(self.ctx, C.byref(n)))
        return n.value

    def LaunchCount(self):
        return _capi.lib().lmn_ctx_launch_count(self.ctx)

    def Close(self):
        for pl in self._plans.values():
            _capi.lib().lmn_plan_destroy(pl)
        self._plans = {}
        for st in list(getattr(self, "_stores", [])):
            st.free()
        if self.ctx:
            _capi.lib().lmn_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass

    # -- reference surface --
    def Field(self):
        return self.ptField

    def MulCounter(self):
        return self.mulCounter

    def SetRingSwitchServer(self, rs):
        self.rs = rs

    def RingSwitch(self):
        return self.rs

    def LoadGaloisKey(self, galEl, evk):
        """rlwe.MemEvaluationKeySet entry: evk = gk.Value[d][0][c] (Q then P limbs), NTT + Montgomery"""
        check(_capi.lib().lmn_load_galois_key(self.ctx, int(galEl), _ptr(_u64(evk))))

    def NewStore(self, count, level=None):
        return CiphertextStore(self, count, level)

    def _plan(self, cols, rhoInv):
        key = (cols, rhoInv)
        if key not in self._plans:
            h = C.c_void_p()
            check(_capi.lib().lmn_plan_create(self.ctx, cols, rhoInv, _ptr(self.ptField.roots), self.ptField.N(), C.byref(h)))
            self._plans[key] = h
        return self._plans[key]

    def PlanCounts(self, cols, rhoInv):
        a, s, m = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(_capi.lib().lmn_plan_counts(self._plan(cols, rhoInv), C.byref(a), C.byref(s), C.byref(m)))
        return a.value, s.value, m.value

    # operator level: operands are (store, index) pairs
    def Add(self, op0, op1, opOut):
        check(_capi.lib().lmn_ct_add(self.ctx, op0[0].h, op0[1], op1[0].h, op1[1], opOut[0].h, opOut[1]))

    def Sub(self, op0, op1, opOut):
        check(_capi.lib().lmn_ct_sub(self.ctx, op0[0].h, op0[1], op1[0].h, op1[1], opOut[0].h, opOut[1]))

    def Mul(self, op0, scalar, opOut):
        self.mulCounter += 1
        check(_capi.lib().lmn_ct_mul_scalar(self.ctx, op0[0].h, op0[1], int(scalar), opOut[0].h, opOut[1]))

    def MulNew(self, op0, plaintext, opOut):
        self.mulCounter += 1
        check(_capi.lib().lmn_ct_mul_pt(self.ctx, op0[0].h, op0[1], _ptr(_u64(plaintext)), opOut[0].h, opOut[1]))

    def InnerSum(self, op, batchSize, n):
        if batchSize != 1:
            raise ValueError("the path uses InnerSum(ct, 1, rows) only (fhe/ligero.go:325)")
        check(_capi.lib().lmn_ct_inner_sum(self.ctx, op[0].h, op[1], n))

    def Rescale(self, op0, opOut):
        check(_capi.lib().lmn_ct_rescale(self.ctx, op0[0].h, op0[1], opOut[0].h, opOut[1]))

    def CopyNew(self, op0, opOut):
        check(_capi.lib().lmn_ct_copy(self.ctx, op0[0].h, op0[1], opOut[0].h, opOut[1]))

    def RescaleScale(self, from_level, to_level):
        return _capi.lib().lmn_rescale_scale(self.ctx, from_level, to_level)

    # -- bgv.Encoder / rlwe.Encryptor halves of ServerBFV (fhe/bfv.go:13-27), on the device --
    def LoadPlaintextRing(self, psi_t=None):
        psi_t = self.params.psi_t if psi_t is None else psi_t
        if psi_t is None:
            raise ValueError("the encoder needs psi_t, the primitive 2N-th root of unity of the plaintext ring")
        check(_capi.lib().lmn_load_plaintext_ring(self.ctx, int(psi_t)))
        self._ringT = True

    def _need_ringT(self):
        if not getattr(self, "_ringT", False):
            self.LoadPlaintextRing()

    def LoadPublicKey(self, pk):
        """pk: [2][k][N] = pk.Value[0].Q, pk.Value[1].Q in NTT + Montgomery form (rlwe.PublicKey as Lattigo holds it)"""
        self._need_ringT()
        check(_capi.lib().lmn_load_public_key(self.ctx, _ptr(_u64(pk))))

    def Encode(self, values):
        """bgv.Encoder.Encode(values, pt) at max level (fhe/ligero.go:204-205): returns the plaintext limbs [k][N]"""
        self._need_ringT()
        v = _u64(values)
        pt = np.empty((len(self.params.Q), self.params.N), np.uint64)
        check(_capi.lib().lmn_encode_plaintext(self.ctx, _ptr(v), len(v), _ptr(pt)))
        return pt

    def EncryptColumns(self, rows, cols, encSeed, firstCol=0, count=None, values=None, witnessKey=None, out=None):
        """"Encrypt matrix" of the demo server (cmd/server/main.go:187-207): EncryptNew(Encode(column)) for `count` columns from
        firstCol.  values: [count][rows] column-major witness values, or None to generate core.RandomMatrixRowMajor's ChaCha20
        witness on the device (witnessKey: 32 bytes, default the reference's LE64(1) key).  Returns the matrix store."""
        count = cols - firstCol if count is None else count
        out = self.NewStore(count) if out is None else out
        key = None
        vals = None
        if values is None:
            key = bytes(witnessKey) if witnessKey is not None else REFERENCE_WITNESS_KEY
        else:
            vals = _u64(values)
            if vals.shape != (count, rows):
                raise ValueError("values must be [count][rows]")
        check(_capi.lib().lmn_encrypt_columns(self.ctx, None if vals is None else _ptr(vals), key, rows, cols, firstCol, count, bytes(encSeed), out.h))
        return out

    def EncryptZero(self, encSeed, streamId=0xFFFFFFFF, out=None, index=0):
        """zeroCol of fhe.Encode (fhe/code.go:15-22)"""
        out = self.NewStore(1) if out is None else out
        check(_capi.lib().lmn_encrypt_zero(self.ctx, bytes(encSeed), streamId, out.h, index))
        return out

    def WitnessColumns(self, rows, cols, firstCol=0, count=None, witnessKey=None):
        """columns of core.RandomMatrixRowMajor's matrix, generated on the device: [count][rows]"""
        count = cols - firstCol if count is None else count
        out = np.empty((count, rows), np.uint64)
        key = bytes(witnessKey) if witnessKey is not None else REFERENCE_WITNESS_KEY
        check(_capi.lib().lmn_witness_columns(self.ctx, key, rows, cols, firstCol, count, _ptr(out)))
        return out


REFERENCE_WITNESS_KEY = (1).to_bytes(8, "little") + bytes(24)     # core/utils.go:55-56: seed = LE64(1) || 0...


class ClientBFV(ServerBFV):
    """ref: fhe/bfv.go:60-106, as far as EncryptedProof.Decrypt uses it: rlwe.Decryptor + bgv.Encoder.Decode on the device
    (SURVEY 8f-4).  sk: [k][N] = sk.Value.Q in NTT + Montgomery form."""

    def __init__(self, ptField, params, sk, device=0, **kw):
        super().__init__(ptField, params, device=device, **kw)
        self.LoadPlaintextRing()
        check(_capi.lib().lmn_load_secret_key(self.ctx, _ptr(_u64(sk))))

    def DecryptDecode(self, store, nValues, first=0, count=None):
        """Decrypt + Decode of store[first : first + count]: the first nValues slots of each, [count][nValues]"""
        count = store.count - first if count is None else count
        out = np.empty((count, nValues), np.uint64)
        check(_capi.lib().lmn_decrypt_decode(self.ctx, store.h, first, count, int(store.scale), nValues, _ptr(out)))
        return out

    def DecryptArrays(self, cts, scale, nValues):
        """the same for ciphertexts that arrived as host arrays [count][2][level+1][N] (EncryptedProof.ReadFrom)"""
        cts = _u64(cts)
        st = CiphertextStore(self, cts.shape[0], cts.shape[2] - 1).upload(cts)
        st.scale = scale
        try:
            return self.DecryptDecode(st, nValues)
        finally:
            st.free()


def NTT(values, size, backend):
    """ref: fhe/ntt.go:12-17 (len(values) == size).  Returns a new store holding the returned slice."""
    if values.count != size:
        raise ValueError("this build evaluates whole-slice transforms: len(values) must equal size")
    out = backend.NewStore(size)
    rho = 1
    check(_capi.lib().lmn_ntt(backend.ctx, backend._plan(size, rho), values.h, out.h))
    a, s, m = backend.PlanCounts(size, rho)
    backend.mulCounter += m
    return out


def Encode(matrix, rows, rhoInv, backend, zeroCol):
    """ref: fhe/code.go:8-34.  zeroCol: (store, index) of the host-produced encryption of the zero vector."""
    cols = matrix.count
    encoded = backend.NewStore(cols * rhoInv)
    check(_capi.lib().lmn_encode(backend.ctx, backend._plan(cols, rhoInv), matrix.h, zeroCol[0].h, zeroCol[1], encoded.h))
    backend.mulCounter += backend.PlanCounts(cols, rhoInv)[2]
    return encoded


def matrixInnerSumEval(matrix, plaintext, rows, backend):
    """ref: fhe/ligero.go:299-370.  plaintext: [k][N] NTT-domain limbs of the bgv plaintext."""
    lout = min(1, backend.params.MaxLevel())
    out = backend.NewStore(matrix.count, lout)
    check(_capi.lib().lmn_inner_product(backend.ctx, matrix.h, _ptr(_u64(plaintext)), rows, out.h))
    out.scale = backend.RescaleScale(backend.params.MaxLevel(), lout)
    if backend.RingSwitch() is not None:
        out = backend.RingSwitch().RingSwitchNew(out)
    return out


class LigeroMetadata:
    def __init__(self, rows, cols, rhoInv, queries):
        self.Rows, self.Cols, self.RhoInv, self.Queries = rows, cols, rhoInv, queries


class EncryptedProof:
    """ref: fhe/ligero.go:185-192; wire format :646-704"""

    def __init__(self, metadata, matR, matZ, queriedCols, merklePaths=None, root=None):
        self.Metadata, self.MatR, self.MatZ, self.QueriedCols = metadata, matR, matZ, queriedCols
        self.MerklePaths, self.Root = merklePaths, root

    def MarshalBinary(self, backend, leafFormat, queryFormat=None, out=None):
        """EncryptedProof.MarshalBinary / WriteTo (ligero.go:646-704): the ciphertext streams are serialised on the GPU and copied
        straight into the result.  leafFormat frames MatR / MatZ, queryFormat (default: the same) the queried columns - they differ
        after a ring switch.  out: optional writable buffer (e.g. pinned memory); returns bytes, or the length written into out."""
        md = _capi.LmnLigeroMetadata(self.Metadata.Rows & 0xFFFFFFFF, self.Metadata.Cols & 0xFFFFFFFF, self.Metadata.RhoInv & 0xFFFFFFFF,
                                     self.Metadata.Queries & 0xFFFFFFFF)
        fm = leafFormat.c_struct()
        fq = queryFormat.c_struct() if queryFormat is not None else None
        paths = b"".join(b"".join(p) for p in self.MerklePaths) if self.MerklePaths else b""
        root = bytes(self.Root) if self.Root else b""
        lib = _capi.lib()
        n = C.c_size_t()
        args = (backend.ctx, C.byref(md), self.MatR.h, self.MatZ.h, self.QueriedCols.h, C.byref(fm), C.byref(fq) if fq is not None else None,
                paths, len(paths), root, len(root))
        check(lib.lmn_proof_marshal(*args, None, 0, C.byref(n)))
        if out is not None:
            dst = np.frombuffer(out, dtype=np.uint8) if not isinstance(out, np.ndarray) else out
            check(lib.lmn_proof_marshal(*args, C.c_void_p(dst.ctypes.data), dst.nbytes, C.byref(n)))
            return n.value
        buf = np.empty(n.value, dtype=np.uint8)
        check(lib.lmn_proof_marshal(*args, C.c_void_p(buf.ctypes.data), buf.nbytes, C.byref(n)))
        return buf.tobytes()

    def Decrypt(self, client):
        """ref: fhe/ligero.go:381-502: queried columns -> `rows` slot values each, MatR / MatZ -> slot 0 of every ciphertext.
        Returns Proof.  The stores are used in place when they live in the client's context, else they travel through the host
        (in the reference the proof arrives over HTTP)."""
        def dec(store, n):
            if store.backend is client:
                return client.DecryptDecode(store, n)
            return client.DecryptArrays(store.download(), store.scale, n)
        if self.MatR.N != client.params.N:
            raise ValueError("ring-switched MatR / MatZ decrypt under the small-ring client (RingSwitch.NewClient), not this one")
        q = dec(self.QueriedCols, self.Metadata.Rows)
        r = dec(self.MatR, 1)[:, 0]
        z = dec(self.MatZ, 1)[:, 0]
        return Proof(self.Metadata, self.Root, r, z, q, self.MerklePaths)

    @staticmethod
    def UnmarshalBinary(data, params, leafFormat, queryFormat=None, matLevel=1, matN=None):
        """EncryptedProof.ReadFrom (ligero.go:706-755), host side (numpy): returns (metadata, MatR, MatZ, QueriedCols as uint64 arrays
        [count][2][level+1][N], paths, root).  Used by the round-trip tests; the verifier is outside this package."""
        data = memoryview(data)
        rows, cols = int.from_bytes(data[0:4], "little"), int.from_bytes(data[4:8], "little")
        rhoInv, queries = data[8], int.from_bytes(data[9:11], "little")
        off = 11
        qf = queryFormat or leafFormat

        def read_cts(count, fmt, level, N):
            nonlocal off
            out = np.empty((count, 2, level + 1, N), dtype=np.uint64)
            for i in range(count):
                if bytes(data[off:off + len(fmt.header)]) != fmt.header:
                    raise ValueError("ciphertext header mismatch")
                off += len(fmt.header)
                for p in range(2):
                    if bytes(data[off:off + len(fmt.polyHeader)]) != fmt.polyHeader:
                        raise ValueError("polynomial header mismatch")
                    off += len(fmt.polyHeader)
                    for l in range(level + 1):
                        if bytes(data[off:off + len(fmt.limbHeader)]) != fmt.limbHeader:
                            raise ValueError("limb header mismatch")
                        off += len(fmt.limbHeader)
                        out[i, p, l] = np.frombuffer(data[off:off + 8 * N], dtype="<u8")
                        off += 8 * N
            return out
        N = 1 << params.logN
        lq = min(1, params.MaxLevel())
        matR = read_cts(cols, leafFormat, matLevel, matN or N)
        matZ = read_cts(cols, leafFormat, matLevel, matN or N)
        qc = read_cts(queries, qf, lq, N)
        merkleLen = cols * rhoInv
        depth = max(0, (merkleLen - 1).bit_length())     # ligero.go:737-739: log2 of the next power of two
        paths = []
        for _ in range(queries):
            paths.append([bytes(data[off + 32 * j:off + 32 * j + 32]) for j in range(depth)])
            off += 32 * depth
        root = bytes(data[off:off + 32])
        return LigeroMetadata(rows, cols, rhoInv, queries), matR, matZ, qc, paths, root


class Proof:
    """ref: fhe/ligero.go:372-379: the decrypted proof (MatR / MatZ as field elements, queried columns as slot values)"""

    def __init__(self, metadata, root, matR, matZ, queriedCols, merklePaths):
        self.Metadata, self.Root, self.MatR, self.MatZ, self.QueriedCols, self.MerklePaths = metadata, root, matR, matZ, queriedCols, merklePaths


class LeafFormat:
    """Framing of the leaf byte stream `ct.WriteTo(buf)` (fhe/ligero.go:157-158): header || 2 x (polyHeader || limbs x
    (limbHeader || N little-endian u64)).  The framing bytes are the Go host's (Lattigo is not vendored in the reference):
    `with_prefixes` puts the u64 length prefixes of lattigo's structs.Vector / structs.Matrix writers around caller-provided
    leading bytes; `placeholder` fills those leading bytes with PLACEHOLDER_METADATA_LEN bytes that are NOT lattigo's MetaData -
    only their LENGTH is pinned, by the serialized sizes the reference's server logs print (tests/test_host_logic.py::
    test_placeholder_framing_reproduces_every_logged_size: every "Marshaled MatR / MatZ / QueriedCols / encrypted proof
    length" line of results/{baseline,experimental}/server/bench_*.txt admits 268..310 MetaData bytes)."""

    PLACEHOLDER_METADATA_LEN = 277

    def __init__(self, header=b"", polyHeader=b"", limbHeader=b""):
        self.header, self.polyHeader, self.limbHeader = bytes(header), bytes(polyHeader), bytes(limbHeader)

    @staticmethod
    def with_prefixes(leading, level, N):
        """leading bytes (the host's MetaData section) || u64(2 polys); per poly u64(level + 1); per limb u64(N)"""
        le = lambda v: int(v).to_bytes(8, "little")
        return LeafFormat(bytes(leading) + le(2), le(level + 1), le(N))

    @staticmethod
    def placeholder(level, N, metadata_len=None):
        """has-metadata flag byte + metadata_len filler bytes + the length prefixes: the right SIZE, not lattigo's bytes"""
        m = LeafFormat.PLACEHOLDER_METADATA_LEN if metadata_len is None else metadata_len
        filler = (b'{"placeholder":"not lattigo MetaData; the Go host supplies the real bytes"}' * 8)[:m].ljust(m, b" ")
        return LeafFormat.with_prefixes(b"\x01" + filler, level, N)

    def stream_len(self, level, N):
        """bytes of one ct.WriteTo stream with this framing"""
        return len(self.header) + 2 * (len(self.polyHeader) + (level + 1) * (len(self.limbHeader) + 8 * N))

    def c_struct(self):
        return _capi.LmnLeafFormat(self.header, len(self.header), self.polyHeader, len(self.polyHeader), self.limbHeader, len(self.limbHeader))


class MerkleTree:
    """ref: core/tree.go:15-21, 76-219.  Built on the GPU (leaf serialisation + SHA-256), node hashes held on the host."""

    def __init__(self, backend, leaves, leafFormat, leafHashes=None):
        """leafHashes: 32*n bytes of leaf hashes gathered from all ranks (multi-GPU Commit) instead of a leaves store"""
        self.h = C.c_void_p()
        if leafHashes is not None:
            check(_capi.lib().lmn_merkle_from_leaf_hashes(backend.ctx, bytes(leafHashes), len(leafHashes) // 32, C.byref(self.h)))
        else:
            fmt = leafFormat.c_struct()
            check(_capi.lib().lmn_merkle_build(backend.ctx, leaves.h, C.byref(fmt), C.byref(self.h)))
        self.depth = _capi.lib().lmn_merkle_depth(self.h)
        self.n = _capi.lib().lmn_merkle_leaf_count(self.h)

    def MerkleRoot(self):
        out = C.create_string_buffer(32)
        check(_capi.lib().lmn_merkle_root(self.h, out))
        return out.raw

    def LeafHash(self, index):
        out = C.create_string_buffer(32)
        check(_capi.lib().lmn_merkle_leaf_hash(self.h, int(index), out))
        return out.raw

    def GetMerklePath(self, index):
        if not 0 <= int(index) < self.n:
            raise IndexError("index %d out of bounds for %d leaves" % (index, self.n))   # tree.go:181-183
        out = C.create_string_buffer(32 * max(1, self.depth))
        check(_capi.lib().lmn_merkle_path(self.h, int(index), out))
        return [out.raw[32 * i:32 * i + 32] for i in range(self.depth)]

    def free(self):
        if self.h:
            _capi.lib().lmn_merkle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def VerifyMerklePath(leafBytes, path, root, index):
    """ref: core/tree.go:223-268 (host side; used by the tests as the reference's own acceptance check)"""
    import hashlib
    cur = hashlib.sha256(leafBytes).digest()
    for sib in path:
        cur = hashlib.sha256(cur + sib if index % 2 == 0 else sib + cur).digest()
        index //= 2
    return cur == root


def LeafHashes(backend, leaves, leafFormat):
    """SHA-256 of every leaf of a (rank-local) leaves store: 32*count bytes"""
    fmt = leafFormat.c_struct()
    out = C.create_string_buffer(32 * leaves.count)
    check(_capi.lib().lmn_leaf_hashes(backend.ctx, leaves.h, C.byref(fmt), out))
    return out.raw


def LeafBytes(backend, leaves, index, leafFormat):
    fmt = leafFormat.c_struct()
    n = C.c_size_t()
    check(_capi.lib().lmn_leaf_bytes(backend.ctx, leaves.h, int(index), C.byref(fmt), None, 0, C.byref(n)))
    buf = C.create_string_buffer(n.value)
    check(_capi.lib().lmn_leaf_bytes(backend.ctx, leaves.h, int(index), C.byref(fmt), buf, n.value, C.byref(n)))
    return buf.raw


class LigeroProver:
    """ref: fhe/ligero.go:32-37, 194-291"""

    def __init__(self, committer, matrix, encoded, leaves=None, tree=None):
        self.Committer, self.Matrix, self.EncodedMatrix = committer, matrix, encoded
        self.Leaves, self.Tree = leaves, tree     # level-1 leaves kept on the device: the query stage becomes a gather
        self.Root = tree.MerkleRoot() if tree is not None else None

    def Prove(self, rPt, bPt, queryIndices, backend):
        """rPt / bPt: plaintext limbs of backend.Encode(r) and backend.Encode(b) (ligero.go:202-221);
        queryIndices: sampleQueryIndices(transcript, Queries, cols*rhoInv) (ligero.go:266, 638-644)."""
        md = self.Committer
        lib = _capi.lib()
        # r / b given as slot values (what transcript.SampleUints and the powers of z^cols produce, ligero.go:202-216) are encoded on
        # the device exactly as backend.Encode does (ligero.go:204-205, 218-219); plaintext limbs [k][N] pass through
        if np.ndim(rPt) == 1:
            rPt = backend.Encode(rPt)
        if np.ndim(bPt) == 1:
            bPt = backend.Encode(bPt)
        lout = min(1, backend.params.MaxLevel())
        matR = backend.NewStore(self.Matrix.count, lout)
        matZ = backend.NewStore(self.Matrix.count, lout)
        # InnerProduct(Matrix, r) and InnerProduct(Matrix, b): the two goroutines of ligero.go:231-242, fused
        check(lib.lmn_inner_product2(backend.ctx, self.Matrix.h, _ptr(_u64(rPt)), _ptr(_u64(bPt)), md.Rows, matR.h, matZ.h))
        matR.scale = matZ.scale = backend.RescaleScale(backend.params.MaxLevel(), lout)
        if backend.RingSwitch() is not None:
            matR = backend.RingSwitch().RingSwitchNew(matR)
            matZ = backend.RingSwitch().RingSwitchNew(matZ)
        # Query columns (ligero.go:262-280)
        idx = np.ascontiguousarray(queryIndices, dtype=np.uint32)
        if len(idx) != md.Queries:
            raise ValueError("expected %d query indices" % md.Queries)
        queried = backend.NewStore(len(idx), lout)
        if self.Leaves is not None:
            # the leaves of Commit are these columns already rescaled (the reference rescales them a second time, ligero.go:269-273)
            check(lib.lmn_store_gather(backend.ctx, self.Leaves.h, _ptr(idx), len(idx), queried.h))
        else:
            check(lib.lmn_query_columns(backend.ctx, self.EncodedMatrix.h, _ptr(idx), len(idx), queried.h))
        queried.scale = backend.RescaleScale(backend.params.MaxLevel(), lout)
        paths = [self.Tree.GetMerklePath(int(i)) for i in idx] if self.Tree is not None else None   # ligero.go:275-279
        return EncryptedProof(LigeroMetadata(md.Rows, md.Cols, md.RhoInv, md.Queries), matR, matZ, queried, paths, self.Root)


class LigeroCommitter(LigeroMetadata):
    """ref: fhe/ligero.go:26-63, 95-124"""

    def __init__(self, securityBits, rows, cols, rhoInv):
        if rows * cols <= 0:
            raise ValueError("size must be positive")
        if securityBits <= 0:
            raise ValueError("securityBits must be positive")
        super().__init__(int(rows), int(cols), int(rhoInv), calculateQueries(securityBits, rhoInv))

    def Commit(self, matrix, backend, zeroCol, leafFormat=None):
        """ref: ligero.go:95-124.  Encode span, then (when a leaf framing is given) the "Merkle tree built" span: leaves =
        encoded columns rescaled to level 1 (processLeafParallel, :126-183), serialised and hashed into core.NewTree.
        Returns the prover; the reference's second return value is prover.Root."""
        if matrix.count != self.Cols:
            raise ValueError("matrix must hold one ciphertext per column")
        encoded = Encode(matrix, self.Rows, self.RhoInv, backend, zeroCol)
        if leafFormat is None:
            return LigeroProver(self, matrix, encoded)
        lout = min(1, backend.params.MaxLevel())
        leaves = backend.NewStore(encoded.count, lout)
        check(_capi.lib().lmn_commit_leaves(backend.ctx, encoded.h, leaves.h))
        leaves.scale = backend.RescaleScale(backend.params.MaxLevel(), lout)
        tree = MerkleTree(backend, leaves, leafFormat)
        return LigeroProver(self, matrix, encoded, leaves, tree)
