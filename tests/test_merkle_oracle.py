"""CPU tests of the Merkle part of the oracle (oracle/merkle_oracle.c) against independent known answers:
FIPS 180-4 SHA-256 vectors, hashlib, and the reference's own acceptance rule VerifyMerklePath (core/tree.go:223-268)."""
import hashlib

import numpy as np
import pytest

from oracle import oracle as O

# FIPS 180-4 / NIST CAVP known answers
KAT = [
    (b"", "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"),
    (b"abc", "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"),
    (b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq", "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"),
    (b"a" * 1000000, "cdc76e5c9914fb9281a1c7e284d73e67f1809a48a497200e046d39ccc7112cd0"),
]


@pytest.mark.parametrize("msg,want", KAT)
def test_sha256_known_answers(msg, want):
    assert O.sha256(msg).hex() == want


def test_sha256_all_padding_lengths():
    rng = np.random.default_rng(3)
    for n in list(range(0, 130)) + [255, 256, 257, 4095, 4096, 65536 + 7]:
        m = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
        assert O.sha256(m) == hashlib.sha256(m).digest(), n


def py_tree(leaf_buffers):
    """core.NewTree restated with hashlib: levels of hashes, an unpaired last node is paired with itself (tree.go:117-147)"""
    lv = [hashlib.sha256(b).digest() for b in leaf_buffers]
    levels = [lv]
    while len(lv) > 1:
        lv = [hashlib.sha256(lv[i] + (lv[i + 1] if i + 1 < len(lv) else lv[i])).digest() for i in range(0, len(lv), 2)]
        levels.append(lv)
    return levels


def verify_path(leaf, path, root, index):
    """core.VerifyMerklePath (tree.go:223-268)"""
    cur = hashlib.sha256(leaf).digest()
    for sib in path:
        cur = hashlib.sha256(cur + sib if index % 2 == 0 else sib + cur).digest()
        index //= 2
    return cur == root


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 7, 8, 9, 31, 32, 33, 100])
def test_tree_and_paths(n):
    rng = np.random.default_rng(n)
    leaves = [rng.integers(0, 256, size=int(rng.integers(1, 300)), dtype=np.uint8).tobytes() for _ in range(n)]
    t = O.MerkleTree(leaves)
    levels = py_tree(leaves)
    assert t.root() == levels[-1][0]
    assert t.total == sum(len(l) for l in levels)
    for i in range(n):
        assert t.leaf_hash(i) == levels[0][i]
        p = t.path(i)
        assert len(p) == len(levels) - 1
        assert verify_path(leaves[i], p, t.root(), i)
        if n > 1:
            assert not verify_path(leaves[i] + b"x", p, t.root(), i)


def test_leaf_bytes_layout():
    rng = np.random.default_rng(5)
    ct = rng.integers(0, 2**58, size=(2, 2, 16), dtype=np.uint64)
    hdr, ph, lh = b"meta!", (2).to_bytes(8, "little"), (16).to_bytes(8, "little")
    want = hdr + b"".join(ph + b"".join(lh + ct[p, l].astype("<u8").tobytes() for l in range(2)) for p in range(2))
    assert O.leaf_bytes(ct, hdr, ph, lh) == want
    assert O.leaf_bytes(ct) == ct.astype("<u8").tobytes()


def test_proof_bytes_layout_and_unmarshal():
    """EncryptedProof.WriteTo (fhe/ligero.go:660-704): metadata || MatR || MatZ || QueriedCols || paths || root, written by the
    oracle, spelled out independently here, and read back by the host-side ReadFrom mirror (fhe.EncryptedProof.UnmarshalBinary)."""
    from lumenos_b200 import fhe
    rng = np.random.default_rng(9)
    rows, cols, rho, nq, N = 70000, 4, 2, 5, 16
    matR = rng.integers(0, 2**58, size=(cols, 2, 2, N), dtype=np.uint64)
    matZ = rng.integers(0, 2**58, size=(cols, 2, 2, N), dtype=np.uint64)
    qc = rng.integers(0, 2**58, size=(nq, 2, 2, N), dtype=np.uint64)
    fmt = (b"meta!", (2).to_bytes(8, "little"), (N).to_bytes(8, "little"))
    depth = 3   # cols * rho = 8 leaves
    paths = [[rng.bytes(32) for _ in range(depth)] for _ in range(nq)]
    root = rng.bytes(32)
    got = O.proof_bytes(rows, cols, rho, nq, matR, matZ, qc, fmt, paths=paths, root=root)
    ser = lambda ct: O.leaf_bytes(ct, *fmt)
    want = (rows.to_bytes(4, "little") + cols.to_bytes(4, "little") + bytes([rho]) + nq.to_bytes(2, "little")
            + b"".join(ser(c) for c in matR) + b"".join(ser(c) for c in matZ) + b"".join(ser(c) for c in qc)
            + b"".join(b"".join(p) for p in paths) + root)
    assert got == want

    class P:   # the two parameters ReadFrom needs
        logN = 4
        @staticmethod
        def MaxLevel():
            return 1
    md, r2, z2, q2, p2, root2 = fhe.EncryptedProof.UnmarshalBinary(got, P, fhe.LeafFormat(*fmt))
    assert (md.Rows, md.Cols, md.RhoInv, md.Queries) == (rows, cols, rho, nq)
    assert (r2 == matR).all() and (z2 == matZ).all() and (q2 == qc).all() and p2 == paths and root2 == root
    # Go's uint8(RhoInv) / uint16(Queries) conversions truncate silently (ligero.go:760-761)
    t = O.proof_bytes(1, 2, 257, 65537, matR[:0], matZ[:0], qc[:0], fmt)
    assert t == (1).to_bytes(4, "little") + (2).to_bytes(4, "little") + b"\x01" + b"\x01\x00"
