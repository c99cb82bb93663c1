"""lmn_merkle_from_leaf_hashes_noalloc: the Merkle tree for hosts that may not allocate (a prover whose HBM is fully committed
before the proof starts, or whose other threads would be stalled by cudaMalloc / cudaFree).  It builds the tree in a 2 MiB device
array that ships inside the library instead of a cudaMalloc'ed one, so the thing to prove is that it returns exactly what the
allocating entry point lmn_merkle_from_leaf_hashes returns - root, every leaf hash, every path - bit for bit (hashes: tolerance
zero, as in all Merkle tests here), including when the shared array still holds the nodes of an earlier, larger tree.

The file is named so that pytest collects it last: `pytest -x` stops at the first failure and this test must not hide the parity
tests.  It makes no assumption about what ran before it (fresh process or the whole suite)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from helpers import Scenario  # noqa: E402


def _node_count(n):
    total = 0
    while True:
        total += n
        if n == 1:
            return total
        n = (n + 1) // 2


def _tree_allocating(lib, check, ctx, hashes, n):
    """root, leaf hashes and paths through the existing entry point (cudaMalloc + lmn_merkle object)"""
    h = C.c_void_p()
    check(lib.lmn_merkle_from_leaf_hashes(ctx, hashes, n, C.byref(h)))
    try:
        depth = lib.lmn_merkle_depth(h)
        root = C.create_string_buffer(32)
        check(lib.lmn_merkle_root(h, root))
        picks = sorted({0, 1 % n, n // 2, n - 1})
        paths = {}
        for i in picks:
            buf = C.create_string_buffer(max(32 * depth, 1))
            check(lib.lmn_merkle_path(h, i, buf))
            paths[i] = buf.raw[:32 * depth]
        return root.raw, depth, paths
    finally:
        lib.lmn_merkle_destroy(h)


def _path_from_nodes(nodes, n, index):
    """what lmn_merkle_path does, read off the level-by-level node array the no-alloc call returns"""
    out, off, cnt, i = b"", 0, n, index
    while cnt > 1:
        sib = i ^ 1
        if sib >= cnt:
            sib = i
        out += nodes[32 * (off + sib):32 * (off + sib + 1)]
        off, cnt, i = off + cnt, (cnt + 1) // 2, i >> 1
    return out


def test_noalloc_merkle_equals_allocating_entry_point_twice_in_a_row():
    from lumenos_b200 import _capi
    lib, check = _capi.lib(), _capi.check
    be = Scenario(6, 32, 16, 3).backend()
    ctx = be.ctx
    rng = np.random.default_rng(2026)
    # two calls back to back on the same array: the maximum size first, then a small odd tree that lands on the stale nodes
    # (an unpaired last node on several levels); then a single leaf (root == the leaf hash, no kernel at all)
    for n in (_capi_max_leaves(), 309, 1):
        hashes = rng.integers(0, 256, size=32 * n, dtype=np.uint8).tobytes()
        want_root, depth, want_paths = _tree_allocating(lib, check, ctx, hashes, n)
        total = _node_count(n)
        root = C.create_string_buffer(32)
        nodes = C.create_string_buffer(32 * total)
        launches0 = lib.lmn_ctx_launch_count(ctx)
        check(lib.lmn_merkle_from_leaf_hashes_noalloc(ctx, hashes, n, root, nodes, 32 * total))
        assert lib.lmn_ctx_launch_count(ctx) - launches0 == depth          # one k_sha256_nodes launch per level above the leaves
        assert root.raw == want_root, "root differs at n = %d" % n
        assert nodes.raw[:32 * n] == hashes                                 # level 0 = the caller's hashes
        assert nodes.raw[-32:] == want_root
        for i, p in want_paths.items():
            assert _path_from_nodes(nodes.raw, n, i) == p, "path %d differs at n = %d" % (i, n)
        # the root alone (nodes_out = NULL) is the same root
        root2 = C.create_string_buffer(32)
        check(lib.lmn_merkle_from_leaf_hashes_noalloc(ctx, hashes, n, root2, None, 0))
        assert root2.raw == want_root
        if n == 309:   # independent of both entry points: hashlib over the same hashes
            lv = [hashes[32 * i:32 * i + 32] for i in range(n)]
            while len(lv) > 1:
                lv = [hashlib.sha256(lv[i] + lv[min(i + 1, len(lv) - 1)]).digest() for i in range(0, len(lv), 2)]
            assert lv[0] == want_root
    be.Close()


def _capi_max_leaves():
    return 32768          # LMN_MERKLE_NOALLOC_MAX_LEAVES, include/lumenos_b200.h


def test_noalloc_merkle_rejects_what_does_not_fit():
    from lumenos_b200 import _capi
    lib = _capi.lib()
    be = Scenario(6, 32, 16, 3).backend()
    n = _capi_max_leaves() + 1
    hashes = bytes(32 * n)
    root = C.create_string_buffer(32)
    assert lib.lmn_merkle_from_leaf_hashes_noalloc(be.ctx, hashes, n, root, None, 0) == -1          # LMN_ERR_INVALID
    msg = lib.lmn_last_error().decode()
    assert "32768" in msg and "lmn_merkle_from_leaf_hashes" in msg, msg
    assert lib.lmn_merkle_from_leaf_hashes_noalloc(be.ctx, hashes, 0, root, None, 0) == -1
    small = C.create_string_buffer(32 * 4)
    assert lib.lmn_merkle_from_leaf_hashes_noalloc(be.ctx, hashes, 4, root, small, 32 * 4) == -1    # 4 leaves have 7 nodes
    assert "nodes_out" in lib.lmn_last_error().decode()
    # a refused call leaves the entry point usable
    assert lib.lmn_merkle_from_leaf_hashes_noalloc(be.ctx, hashes, 4, root, None, 0) == 0
    be.Close()
