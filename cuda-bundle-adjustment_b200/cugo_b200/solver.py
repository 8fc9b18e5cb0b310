"""ctypes binding of lib/libcugo_b200.so (C ABI: include/cugo_b200.h).

`BlockSolver` mirrors the reference's cugo::BlockSolver method names
(src/block_solver.h:54-173) and `optimize` mirrors
CudaGraphOptimisationImpl::optimize (src/cuda_graph_optimisation.cpp:48-154).
There is no CPU fallback: if the library or a CUDA device is missing, construction
raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .pack import Problem

_HERE = os.path.dirname(os.path.abspath(__file__))
# CUGO_B200_LIB: another build of the same library (A/B measurements of kernel variants: scripts/ab_lib.sh)
LIB_PATH = os.environ.get("CUGO_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libcugo_b200.so")

RK = dict(none=0, cauchy=1, tukey=2, huber=3)
STRUCT = dict(hpl_colptr=0, hpl_rowind=1, hsc_rowptr=2, hsc_colind=3, csr_rowptr=4, csr_colind=5,
              bsr2csr=6, mulids=7, edge2hpl=8, perm=9, sn_start=10, sn_rowptr=11, sn_rows=12, sn_parent=13, sn_level=14,
              level_sn=15, child_idx=16, inv_map=17, asm_blk=18, asm_row=19, asm_col=20)
VALUES = dict(Hpp=0, Hll=1, Hpl=2, b=3, x=4, bsc=5, Hsc=6, invHll=7, chi_per_edge=8)
SIZES = ("P", "L", "E", "nHpl", "nblk", "nmul", "nnzSymm", "nsupernodes", "nlevels", "kernels_per_trial")

# every symbol include/cugo_b200.h declares
SYMBOLS = (
    "cugo_b200_create", "cugo_b200_destroy", "cugo_b200_last_error", "cugo_b200_set_vertices",
    "cugo_b200_clear_edge_sets", "cugo_b200_add_edge_set", "cugo_b200_set_robust_kernel",
    "cugo_b200_set_outliers", "cugo_b200_set_estimates", "cugo_b200_build_structure",
    "cugo_b200_compute_errors", "cugo_b200_build_system", "cugo_b200_max_diagonal",
    "cugo_b200_set_lambda", "cugo_b200_restore_diagonal", "cugo_b200_solve", "cugo_b200_update",
    "cugo_b200_compute_scale", "cugo_b200_push", "cugo_b200_pop", "cugo_b200_optimize",
    "cugo_b200_get_estimates", "cugo_b200_get_structure", "cugo_b200_get_values", "cugo_b200_get_sizes",
    "cugo_b200_get_last_timing", "cugo_b200_stream", "cugo_b200_profile_trial", "cugo_b200_update_measurements",
    "cugo_b200_debug_counters", "cugo_b200_get_factor_stats", "cugo_b200_comm_unique_id", "cugo_b200_comm_init",
    "cugo_b200_chi2_noalloc", "cugo_b200_set_outlier_threshold", "cugo_b200_update_edges", "cugo_b200_get_outliers",
    "cugo_b200_set_option", "cugo_b200_add_depth_edge_set", "cugo_b200_add_plane_edge_set",
    "cugo_b200_add_line_edge_set",
)

_lib = None


class CugoError(RuntimeError):
    pass


def load_library():
    """dlopen the product library (does not touch the GPU)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CugoError("%s not found: run `make -C cuda-bundle-adjustment_b200` (or "
                            "__graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        lib.cugo_b200_last_error.restype = C.c_char_p
        lib.cugo_b200_stream.restype = C.c_void_p
        _lib = lib
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class BlockSolver:
    def __init__(self, device=0, f32=False):
        self.lib = load_library()
        self.h = C.c_void_p()
        rc = self.lib.cugo_b200_create(int(device), 1 if f32 else 0, C.byref(self.h))
        if rc != 0:
            raise CugoError("cugo_b200_create failed with %d (%s)" % (
                rc, {-2: "no CUDA device; this engine has no CPU path", -3: "CUDA error"}.get(rc, "")))
        self.prob = None

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.cugo_b200_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise CugoError("cugo_b200 error %d: %s" % (rc, self.lib.cugo_b200_last_error(self.h).decode()))

    # -- BlockSolver::initialize ------------------------------------------------
    def initialize(self, prob: Problem, robust="none", delta=1.0):
        """Upload vertices and edge sets (robust/delta: one value or a per-set list)."""
        self.prob = prob
        poses, lms = _f64(prob.poses), _f64(prob.landmarks)
        self._ck(self.lib.cugo_b200_set_vertices(self.h, poses.shape[0], prob.nP_act, _p(poses, C.c_double),
                                                 lms.shape[0], prob.nL_act, _p(lms, C.c_double)))
        self._ck(self.lib.cugo_b200_clear_edge_sets(self.h))
        self.set_ids = []
        for i, s in enumerate(prob.sets):
            rk = robust[i] if isinstance(robust, (list, tuple)) else robust
            dl = delta[i] if isinstance(delta, (list, tuple)) else delta
            rkc = RK[rk] if isinstance(rk, str) else int(rk)
            sid = C.c_int()
            kind = getattr(s, "kind", "")
            iP, meas, om = _i32(s.iP), _f64(s.meas), _f64(s.omega)
            if kind in ("plane", "line"):      # PlaneEdgeSet / LineEdgeSet (include/icp_types.h:19-107)
                fn = self.lib.cugo_b200_add_plane_edge_set if kind == "plane" else self.lib.cugo_b200_add_line_edge_set
                self._ck(fn(self.h, s.n, _p(iP, C.c_int), _p(meas, C.c_double), _p(om, C.c_double), int(om.shape[0]),
                            rkc, C.c_double(dl), C.byref(sid)))
            else:
                iL, cam = _i32(s.iL), _f64(s.cam)
                dim = 4 if kind == "depth" else s.dim      # DepthEdgeSet (include/ba_types.h:171-234)
                self._ck(self.lib.cugo_b200_add_edge_set(
                    self.h, dim, s.n, _p(iP, C.c_int), _p(iL, C.c_int), _p(meas, C.c_double),
                    _p(om, C.c_double), int(om.shape[0]), _p(cam, C.c_double), int(cam.shape[0]),
                    rkc, C.c_double(dl), C.byref(sid)))
            self.set_ids.append(sid.value)

    def set_estimates(self, poses, lms):
        poses, lms = _f64(poses), _f64(lms)
        self._ck(self.lib.cugo_b200_set_estimates(self.h, _p(poses, C.c_double), _p(lms, C.c_double)))

    def set_robust_kernel(self, set_idx, robust, delta):
        self._ck(self.lib.cugo_b200_set_robust_kernel(self.h, self.set_ids[set_idx], RK[robust], C.c_double(delta)))

    def set_outliers(self, set_idx, mask):
        m = _i32(mask)
        self._ck(self.lib.cugo_b200_set_outliers(self.h, self.set_ids[set_idx], _p(m, C.c_int)))

    # -- outlier gating (EdgeSet::setOutlierThreshold / updateEdges / getOutlierCount) -------------------
    def set_outlier_threshold(self, set_idx, thr):
        self._ck(self.lib.cugo_b200_set_outlier_threshold(self.h, self.set_ids[set_idx], C.c_double(thr)))

    def update_edges(self):
        self._ck(self.lib.cugo_b200_update_edges(self.h))

    def outliers(self, set_idx):
        """(mask [n] int32 in the set's edge order, outlier count)"""
        m = np.zeros(self.prob.sets[set_idx].n, np.int32)
        no, ni = C.c_int(), C.c_int()
        self._ck(self.lib.cugo_b200_get_outliers(self.h, self.set_ids[set_idx], _p(m, C.c_int), C.byref(no), C.byref(ni)))
        return m, no.value

    # -- stage API ----------------------------------------------------------------
    def build_structure(self):
        self._ck(self.lib.cugo_b200_build_structure(self.h))

    def compute_errors(self):
        v = C.c_double()
        self._ck(self.lib.cugo_b200_compute_errors(self.h, C.byref(v)))
        return v.value

    def build_system(self):
        self._ck(self.lib.cugo_b200_build_system(self.h))

    def max_diagonal(self):
        v = C.c_double()
        self._ck(self.lib.cugo_b200_max_diagonal(self.h, C.byref(v)))
        return v.value

    def set_lambda(self, lam):
        self._ck(self.lib.cugo_b200_set_lambda(self.h, C.c_double(lam)))

    def restore_diagonal(self):
        self._ck(self.lib.cugo_b200_restore_diagonal(self.h))

    def solve(self):
        ok = C.c_int()
        self._ck(self.lib.cugo_b200_solve(self.h, C.byref(ok)))
        return bool(ok.value)

    def update(self):
        self._ck(self.lib.cugo_b200_update(self.h))

    def compute_scale(self, lam):
        v = C.c_double()
        self._ck(self.lib.cugo_b200_compute_scale(self.h, C.c_double(lam), C.byref(v)))
        return v.value

    def push(self):
        self._ck(self.lib.cugo_b200_push(self.h))

    def pop(self):
        self._ck(self.lib.cugo_b200_pop(self.h))

    # -- device-resident LM ---------------------------------------------------------
    def optimize(self, niter):
        cap = max(int(niter), 1)
        chi2 = np.zeros(cap, np.float64)
        lam = np.zeros(cap, np.float64)
        trials = np.zeros(cap, np.int32)
        n = C.c_int()
        self._ck(self.lib.cugo_b200_optimize(self.h, int(niter), cap, _p(chi2, C.c_double), _p(lam, C.c_double),
                                             _p(trials, C.c_int), C.byref(n)))
        k = min(n.value, cap)
        return chi2[:k].copy(), lam[:k].copy(), trials[:k].copy()

    def estimates(self, out=None):
        """(poses [P,7], landmarks [L,3]); `out` = a pair of caller-owned float64 arrays of those shapes (pinned
        host memory makes the read-back an asynchronous DMA)."""
        P, L = self.prob.poses.shape[0], self.prob.landmarks.shape[0]
        if out is not None:
            poses, lms = out
            assert poses.shape == (P, 7) and lms.shape == (L, 3) and poses.dtype == np.float64 and lms.dtype == np.float64
            assert poses.flags.c_contiguous and lms.flags.c_contiguous
        else:
            poses = np.empty((P, 7), np.float64)
            lms = np.empty((L, 3), np.float64)
        self._ck(self.lib.cugo_b200_get_estimates(self.h, _p(poses, C.c_double), _p(lms, C.c_double)))
        return poses, lms

    # -- introspection ----------------------------------------------------------------
    def structure(self, name):
        n = C.c_long()
        self._ck(self.lib.cugo_b200_get_structure(self.h, STRUCT[name], None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 0), np.int32)
        if n.value > 0:
            self._ck(self.lib.cugo_b200_get_structure(self.h, STRUCT[name], _p(out, C.c_int), n.value, C.byref(n)))
        return out

    def values(self, name):
        n = C.c_long()
        self._ck(self.lib.cugo_b200_get_values(self.h, VALUES[name], None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 0), np.float64)
        if n.value > 0:
            self._ck(self.lib.cugo_b200_get_values(self.h, VALUES[name], _p(out, C.c_double), n.value, C.byref(n)))
        return out

    def sizes(self):
        a = (C.c_long * 10)()
        self._ck(self.lib.cugo_b200_get_sizes(self.h, a))
        return dict(zip(SIZES, (int(v) for v in a)))

    def last_timing(self):
        ms, k = C.c_double(), C.c_long()
        self._ck(self.lib.cugo_b200_get_last_timing(self.h, C.byref(ms), C.byref(k)))
        return ms.value, k.value

    def profile_trial(self):
        """Per-kernel device times of one LM trial: list of (kernel name, ms)."""
        cap, ln = 512, 32
        names = C.create_string_buffer(cap * ln)
        ms = np.zeros(cap, np.float64)
        n = C.c_int()
        self._ck(self.lib.cugo_b200_profile_trial(self.h, cap, names, ln, _p(ms, C.c_double), C.byref(n)))
        raw = names.raw
        return [(raw[i * ln:(i + 1) * ln].split(b"\0")[0].decode(), float(ms[i])) for i in range(n.value)]

    def set_option(self, name, value):
        """Scheduling variants of the trial (include/cugo_b200.h: cugo_b200_set_option)."""
        self._ck(self.lib.cugo_b200_set_option(self.h, name.encode(), int(value)))

    def update_measurements(self, set_idx, meas, omega):
        meas, omega = _f64(meas), _f64(omega)
        self._ck(self.lib.cugo_b200_update_measurements(self.h, self.set_ids[set_idx], _p(meas, C.c_double), _p(omega, C.c_double),
                                                        int(omega.shape[0])))

    # -- multi-GPU plumbing (see cugo_b200/dist.py) ---------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        rc = load_library().cugo_b200_comm_unique_id(buf)
        if rc != 0:
            raise CugoError("cugo_b200_comm_unique_id failed with %d (NCCL not loadable?)" % rc)
        return buf.raw

    def comm_init(self, nranks, rank, uid: bytes):
        self._ck(self.lib.cugo_b200_comm_init(self.h, int(nranks), int(rank), C.c_char_p(uid)))

    def factor_stats(self):
        a = (C.c_double * 4)()
        self._ck(self.lib.cugo_b200_get_factor_stats(self.h, a))
        return dict(flops_panel=a[0], flops_update=a[1], flops_backward=a[2], l_doubles=a[3])

    def debug_counters(self, n=64 * 24):
        a = (C.c_longlong * n)()
        self._ck(self.lib.cugo_b200_debug_counters(a, n))
        v = np.array(a[:], dtype=np.int64)
        return v.reshape(-1, 24) if n % 24 == 0 else v

    def stream(self):
        return self.lib.cugo_b200_stream(self.h)
