"""ctypes binding of oracle/ba_oracle.c -- TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

STRUCT = dict(hpl_colptr=0, hpl_rowind=1, hsc_rowptr=2, hsc_colind=3, csr_rowptr=4,
              csr_colind=5, bsr2csr=6, mulids=7, edge2hpl=8)
VALUES = dict(Hpp=0, Hll=1, Hpl=2, b=3, x=4, bsc=5, Hsc=6, invHll=7, poses=8, landmarks=9)
RK = dict(none=0, cauchy=1, tukey=2, huber=3)


def build(force=False):
    """Compile the C restatement (double and float variants)."""
    src = os.path.join(HERE, "ba_oracle.c")
    for name, extra in (("libba_oracle.so", []), ("libba_oracle_f32.so", ["-DBAO_FLOAT"])):
        out = os.path.join(HERE, name)
        if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", *extra,
                                   "-o", out, src, "-lm"])


_libs = {}


def _lib(f32=False):
    key = bool(f32)
    if key not in _libs:
        build()
        lib = C.CDLL(os.path.join(HERE, "libba_oracle_f32.so" if f32 else "libba_oracle.so"))
        lib.bao_create.restype = C.c_void_p
        lib.bao_compute_errors.restype = C.c_double
        lib.bao_max_diagonal.restype = C.c_double
        lib.bao_compute_scale.restype = C.c_double
        lib.bao_structure_size.restype = C.c_long
        lib.bao_values_size.restype = C.c_long
        _libs[key] = lib
    return _libs[key]


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class Oracle:
    def __init__(self, prob, robust="none", delta=1.0, f32=False, threads=1):
        """prob: cugo_b200.pack.Problem.  robust/delta may be a single value or a
        per-set list."""
        self.lib = _lib(f32)
        self.prob = prob
        self.h = C.c_void_p(self.lib.bao_create())
        self.lib.bao_set_threads(self.h, int(threads))
        poses = np.ascontiguousarray(prob.poses, np.float64)
        lms = np.ascontiguousarray(prob.landmarks, np.float64)
        rc = self.lib.bao_set_vertices(self.h, poses.shape[0], prob.nP_act, _p(poses, C.c_double),
                                       lms.shape[0], prob.nL_act, _p(lms, C.c_double))
        assert rc == 0, rc
        for i, s in enumerate(prob.sets):
            rk = robust[i] if isinstance(robust, (list, tuple)) else robust
            dl = delta[i] if isinstance(delta, (list, tuple)) else delta
            rkc = RK[rk] if isinstance(rk, str) else int(rk)
            kind = getattr(s, "kind", "")
            if kind in ("plane", "line"):
                rc = self.lib.bao_add_icp_edge_set(
                    self.h, 2 if kind == "plane" else 3, s.n, _p(s.iP, C.c_int), _p(s.meas, C.c_double),
                    _p(s.omega, C.c_double), int(s.omega.shape[0]), rkc, C.c_double(dl))
            elif kind == "depth":
                rc = self.lib.bao_add_depth_edge_set(
                    self.h, s.n, _p(s.iP, C.c_int), _p(s.iL, C.c_int), _p(s.meas, C.c_double),
                    _p(s.omega, C.c_double), int(s.omega.shape[0]), _p(s.cam, C.c_double),
                    int(s.cam.shape[0]), rkc, C.c_double(dl))
            else:
                rc = self.lib.bao_add_edge_set(
                    self.h, s.dim, s.n, _p(s.iP, C.c_int), _p(s.iL, C.c_int), _p(s.meas, C.c_double),
                    _p(s.omega, C.c_double), int(s.omega.shape[0]), _p(s.cam, C.c_double),
                    int(s.cam.shape[0]), rkc, C.c_double(dl))
            assert rc == 0, rc

    def __del__(self):
        try:
            self.lib.bao_destroy(self.h)
        except Exception:
            pass

    # staged API (names follow the reference BlockSolver, src/block_solver.h:54-173)
    def build_structure(self):
        assert self.lib.bao_build_structure(self.h) == 0

    def compute_errors(self):
        return float(self.lib.bao_compute_errors(self.h))

    def build_system(self):
        assert self.lib.bao_build_system(self.h) == 0

    def max_diagonal(self):
        return float(self.lib.bao_max_diagonal(self.h))

    def set_lambda(self, lam):
        self.lib.bao_set_lambda(self.h, C.c_double(lam))

    def restore_diagonal(self):
        self.lib.bao_restore_diagonal(self.h)

    def solve(self):
        return bool(self.lib.bao_solve(self.h))

    def update(self):
        self.lib.bao_update(self.h)

    def compute_scale(self, lam):
        return float(self.lib.bao_compute_scale(self.h, C.c_double(lam)))

    def push(self):
        self.lib.bao_push(self.h)

    def pop(self):
        self.lib.bao_pop(self.h)

    def set_outliers(self, set_idx, mask):
        m = np.ascontiguousarray(mask, np.int32)
        assert self.lib.bao_set_outliers(self.h, set_idx, _p(m, C.c_int)) == 0

    # outlier gating (EdgeSet::setOutlierThreshold / updateEdges / getOutlierCount, src/optimisable_graph.hpp:415,603-670)
    def set_outlier_threshold(self, set_idx, thr):
        assert self.lib.bao_set_outlier_threshold(self.h, set_idx, C.c_double(thr)) == 0

    def update_edges(self):
        return bool(self.lib.bao_update_edges(self.h))

    def outliers(self, set_idx):
        """(mask [n] int32 in the set's edge order, outlier count)"""
        m = np.zeros(self.prob.sets[set_idx].n, np.int32)
        c = C.c_int()
        assert self.lib.bao_get_outliers(self.h, set_idx, _p(m, C.c_int), C.byref(c)) == 0
        return m, c.value

    def optimize(self, niter):
        chi2 = np.zeros(niter, np.float64)
        lam = np.zeros(niter, np.float64)
        trials = np.zeros(niter, np.int32)
        n = self.lib.bao_optimize(self.h, niter, _p(chi2, C.c_double), _p(lam, C.c_double),
                                  _p(trials, C.c_int))
        return chi2[:n].copy(), lam[:n].copy(), trials[:n].copy()

    def structure(self, name):
        w = STRUCT[name]
        n = self.lib.bao_structure_size(self.h, w)
        out = np.zeros(max(n, 0), np.int32)
        assert self.lib.bao_get_structure(self.h, w, _p(out, C.c_int)) == 0
        return out

    def values(self, name):
        w = VALUES[name]
        n = self.lib.bao_values_size(self.h, w)
        out = np.zeros(max(n, 0), np.float64)
        assert self.lib.bao_get_values(self.h, w, _p(out, C.c_double)) == 0
        return out

    def edge_errors(self, set_idx):
        s = self.prob.sets[set_idx]
        err = np.zeros((s.n, s.dim), np.float64)
        chi = np.zeros(s.n, np.float64)
        assert self.lib.bao_get_edge_errors(self.h, set_idx, _p(err, C.c_double), _p(chi, C.c_double)) == 0
        return err, chi

    def estimates(self):
        return (self.values("poses").reshape(-1, 7), self.values("landmarks").reshape(-1, 3))
