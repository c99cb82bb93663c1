"""CPU tests of the product's host logic: the C-ABI library loads and exports every declared symbol, the plan
compiler reproduces core.NTT (the reference's plaintext twin, via the oracle) and the reference's op counts, and
the Python mirror keeps the reference's argument checks.  No compute call touches a GPU here."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

from lumenos_b200 import _capi, fhe
from lumenos_b200 import build as lb
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KAT = json.load(open(os.path.join(ROOT, "tests", "golden", "kat.json")))
T = KAT["t"]


@pytest.fixture(scope="module")
def lib():
    lb.build()
    return _capi.lib()


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "lumenos_b200.h")).read()
    declared = set(re.findall(r"\b(lmn_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(lib, name), "liblumenos_b200.so does not export %s" % name
    # and the ctypes table covers the header exactly
    assert declared == set(_capi.SIGNATURES), declared ^ set(_capi.SIGNATURES)


def test_missing_library_is_loud(monkeypatch, tmp_path):
    monkeypatch.setattr(_capi, "_lib", None)
    monkeypatch.setattr(_capi, "SO_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _capi.lib()


def _plan_apply(lib, n, roots, values):
    v = np.ascontiguousarray(values, np.uint64).copy()
    counts = np.zeros(3, np.uint64)
    rc = lib.lmn_plan_host_check(n, roots.ctypes.data_as(C.c_void_p), T, v.ctypes.data_as(C.c_void_p),
                                 counts.ctypes.data_as(C.c_void_p))
    assert rc == 0, lib.lmn_last_error()
    return v, [int(x) for x in counts]


@pytest.mark.parametrize("n", [16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192])
def test_plan_reproduces_core_ntt(lib, n):
    roots = O.field_roots_forward(n)
    rng = np.random.default_rng(n)
    vals = rng.integers(0, T, size=n, dtype=np.uint64)
    got, counts = _plan_apply(lib, n, roots, vals)
    assert (got == O.core_ntt(vals, n, roots)).all()
    if str(n) in KAT["ntt_op_counts"]:
        assert counts == KAT["ntt_op_counts"][str(n)]


def test_plan_rejects_bad_sizes(lib):
    roots = O.field_roots_forward(16)
    v = np.zeros(16, np.uint64)
    assert lib.lmn_plan_host_check(12, roots.ctypes.data_as(C.c_void_p), T, v.ctypes.data_as(C.c_void_p), None) != 0
    assert b"power of two" in lib.lmn_last_error()


def test_generate_bgv_params_heuristic():
    lit = fhe.GenerateBGVParamsForNTT(1024, 12, T)
    assert lit["LogQ"] == [58] + [56] * 9 and lit["LogP"] == [55, 55]
    assert len(fhe.GenerateBGVParamsForNTT(4096, 14, T)["LogQ"]) == 12
    with pytest.raises(ValueError, match="does not satisfy T = 1"):
        fhe.GenerateBGVParamsForNTT(1024, 12, 65537 * 3)
    with pytest.raises(ValueError):
        fhe.GenerateBGVParamsForNTT(1, 12, T)


def test_calculate_queries_and_committer():
    assert fhe.calculateQueries(128, 2) == KAT["queries"]
    c = fhe.LigeroCommitter(128, 2048, 1024, 2)
    assert (c.Rows, c.Cols, c.RhoInv, c.Queries) == (2048, 1024, 2, 309)
    with pytest.raises(ValueError):
        fhe.LigeroCommitter(0, 2048, 1024, 2)


def test_galois_elements_for_innersum_match_oracle():
    for logN, rows in [(6, 32), (6, 64), (8, 64), (12, 2048), (12, 4096)]:
        P = O.Params(logN, k=2)
        params = fhe.BGVParameters(logN, P.q, P.p, T, P.psi_q, P.psi_p)
        assert params.GaloisElementsForInnerSum(1, rows) == P.galois_elements_for_innersum(rows)


def test_context_creation_without_gpu_reports_error(lib):
    """No GPU in the CPU test environment: the call must fail with a status and message, not crash or fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    P = O.Params(6, k=2)
    params = fhe.BGVParameters(6, P.q, P.p, T, P.psi_q, P.psi_p)
    with pytest.raises(fhe.LumenosError):
        fhe.ServerBFV(fhe.PrimeField(T, O.field_roots_forward(16)), params)


def test_hostutil_matches_oracle():
    """the Python stand-in for the Go host's Lattigo-derived material agrees with the oracle's restatement"""
    from lumenos_b200 import hostutil as H
    for logN, k in [(6, 3), (12, 10), (14, 12)]:
        P = O.Params(logN, k=k)
        q, p = H.generate_moduli(logN, k)
        assert q == [int(x) for x in P.q] and p == [int(x) for x in P.p]
        assert [H.psi_of(m, logN) for m in q[:2] + p] == [int(x) for x in list(P.psi_q[:2]) + list(P.psi_p)]
    for n in [16, 2048]:
        assert (H.field_roots_forward(n) == O.field_roots_forward(n)).all()
    assert H.num_q_primes_for_ntt(4096) == 12
