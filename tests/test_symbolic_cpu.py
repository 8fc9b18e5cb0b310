"""CPU tests of csrc/symbolic.h (the host/device-shared symbolic analysis) through the
host harness tests/csrc/mf_host_test.cpp: ordering is a permutation, the multifrontal data flow
the CUDA kernels implement solves H x = b to rounding on random block patterns."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def lib():
    src = os.path.join(HERE, "csrc", "mf_host_test.cpp")
    out = os.path.join(HERE, "csrc", "libmf_host_test.so")
    hdr = os.path.join(HERE, "..", "cuda-bundle-adjustment_b200", "csrc", "symbolic.h")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", out, src])
    return C.CDLL(out)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def random_spd_blocks(P, extra_edges, rng, chain=True, components=1):
    """Upper BSR pattern (diag + chain + random blocks) and SPD values."""
    pat = np.eye(P, dtype=bool)
    comp = np.arange(P) * components // max(P, 1)
    if chain:
        for i in range(P - 1):
            if comp[i] == comp[i + 1]:
                pat[i, i + 1] = True
    for _ in range(extra_edges):
        i, j = sorted(rng.integers(0, P, 2))
        if comp[i] == comp[j]:
            pat[i, j] = True
    rowptr = np.concatenate([[0], np.cumsum(pat.sum(1))]).astype(np.int32)
    colind = np.concatenate([np.flatnonzero(pat[r]) for r in range(P)]).astype(np.int32)
    n = 6 * P
    A = np.zeros((n, n))
    vals = np.zeros((len(colind), 36))
    k = 0
    for r in range(P):
        for c in colind[rowptr[r]:rowptr[r + 1]]:
            B = rng.normal(size=(6, 6))
            if r == c:
                B = B @ B.T
            A[6 * r:6 * r + 6, 6 * c:6 * c + 6] = B
            A[6 * c:6 * c + 6, 6 * r:6 * r + 6] = B.T
            k += 1
    A += np.eye(n) * (np.abs(A).sum(1).max() + 1.0)       # diagonally dominant => SPD
    k = 0
    for r in range(P):
        for c in colind[rowptr[r]:rowptr[r + 1]]:
            vals[k] = A[6 * r:6 * r + 6, 6 * c:6 * c + 6].T.reshape(-1)   # column-major block
            k += 1
    return rowptr, colind, vals.reshape(-1), A


@pytest.mark.parametrize("P,extra,leaf,maxw,comps", [
    (1, 0, 16, 16, 1), (2, 0, 16, 16, 1), (7, 5, 2, 1, 1), (30, 20, 4, 2, 1), (64, 40, 8, 4, 1),
    (120, 150, 16, 16, 1), (90, 60, 8, 8, 3), (50, 2000, 8, 8, 1), (40, 0, 4, 4, 40)])
def test_multifrontal_solves(lib, P, extra, leaf, maxw, comps):
    rng = np.random.default_rng(P * 1000 + extra)
    rp, ci, vals, A = random_spd_blocks(P, extra, rng, components=comps)
    rhs = rng.normal(size=6 * P)
    x = np.zeros(6 * P); perm = np.zeros(P, np.int32); cnt = np.zeros(6, np.int32); sz = np.zeros(4)
    ok = lib.cugo_sym_test(P, _p(rp, C.c_int), _p(ci, C.c_int), leaf, maxw, _p(vals, C.c_double),
                           _p(rhs, C.c_double), _p(x, C.c_double), _p(perm, C.c_int), _p(cnt, C.c_int),
                           _p(sz, C.c_double))
    assert ok == 1 and cnt[5] == 0
    assert sorted(perm.tolist()) == list(range(P))
    assert cnt[4] <= maxw
    xref = np.linalg.solve(A, rhs)
    np.testing.assert_allclose(x, xref, rtol=1e-9, atol=1e-11 * np.abs(xref).max())


def test_non_positive_pivot_reported(lib):
    rng = np.random.default_rng(3)
    rp, ci, vals, A = random_spd_blocks(10, 5, rng)
    vals = vals.copy()
    vals[:36] = 0.0        # zero the first diagonal block
    rhs = np.ones(60); x = np.zeros(60); perm = np.zeros(10, np.int32); cnt = np.zeros(6, np.int32); sz = np.zeros(4)
    ok = lib.cugo_sym_test(10, _p(rp, C.c_int), _p(ci, C.c_int), 4, 4, _p(vals, C.c_double), _p(rhs, C.c_double),
                           _p(x, C.c_double), _p(perm, C.c_int), _p(cnt, C.c_int), _p(sz, C.c_double))
    assert ok == 0
