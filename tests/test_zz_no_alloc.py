"""cugo_b200_chi2_noalloc (include/cugo_b200.h): the allocation-free chi2 entry point, called through the C ABI
with caller-owned DEVICE buffers, against the existing cugo_b200_compute_errors on the same graph.  Collected
last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CHI2_RTOL = 1e-12        # same tolerance as test_stages_match_oracle's chi2 check (summation order only)
MAX_EDGES = 8 * 1024 * 1024


def _noalloc_chi2(lib, prob, poses, lms, robust, delta, stream=None):
    """Sum over the problem's edge sets of the entry point's result; all buffers are torch CUDA tensors."""
    import torch
    dev = torch.device("cuda:0")
    d_poses = torch.from_numpy(np.ascontiguousarray(poses, np.float64)).to(dev)
    d_lms = torch.from_numpy(np.ascontiguousarray(lms, np.float64)).to(dev)
    total = 0.0
    for s in prob.sets:
        d_iP = torch.from_numpy(s.iP).to(dev)
        d_iL = torch.from_numpy(s.iL).to(dev)
        d_meas = torch.from_numpy(s.meas).to(dev)
        d_om = torch.from_numpy(s.omega).to(dev)
        d_out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
        cam = (C.c_double * 5)(*[float(v) for v in s.cam[0]])
        err = C.create_string_buffer(256)
        torch.cuda.synchronize()
        rc = lib.cugo_b200_chi2_noalloc(0, C.c_void_p(stream), s.dim, s.n, C.c_void_p(d_iP.data_ptr()),
                                        C.c_void_p(d_iL.data_ptr()), C.c_void_p(d_meas.data_ptr()),
                                        C.c_void_p(d_om.data_ptr()), int(s.omega.shape[0]), cam, robust,
                                        C.c_double(delta), C.c_void_p(d_poses.data_ptr()), C.c_void_p(d_lms.data_ptr()),
                                        C.c_void_p(d_out.data_ptr()), err, 256)
        assert rc == 0, err.value.decode()
        torch.cuda.synchronize()
        total += float(d_out.cpu()[0])
    return total


@pytest.mark.parametrize("robust,delta", [("none", 1.0), ("huber", 2.4)])
def test_noalloc_chi2_matches_compute_errors_twice(graphs, robust, delta):
    from cugo_b200 import solver
    g, prob = graphs("small")
    lib = solver.load_library()
    s = solver.BlockSolver(0)
    s.initialize(prob, robust=robust, delta=delta)
    rk = solver.RK[robust]
    # call 1: the original estimates
    F1 = s.compute_errors()
    G1 = _noalloc_chi2(lib, prob, prob.poses, prob.landmarks, rk, delta)
    assert G1 == pytest.approx(F1, rel=CHI2_RTOL)
    # call 2, straight after: other estimates (the static scratch still holds call 1's partial sums)
    rng = np.random.default_rng(5)
    lms2 = prob.landmarks + 0.05 * rng.standard_normal(prob.landmarks.shape)
    s.set_estimates(prob.poses, lms2)
    F2 = s.compute_errors()
    G2 = _noalloc_chi2(lib, prob, prob.poses, lms2, rk, delta)
    assert F2 != F1
    assert G2 == pytest.approx(F2, rel=CHI2_RTOL)
    # and the first inputs again: bit-identical to the first answer (fixed summation order)
    assert _noalloc_chi2(lib, prob, prob.poses, prob.landmarks, rk, delta) == G1
    s.close()


def test_noalloc_rejects_oversize_and_bad_arguments():
    from cugo_b200 import solver
    lib = solver.load_library()
    err = C.create_string_buffer(256)
    cam = (C.c_double * 5)(1, 1, 0, 0, 0)
    one = C.c_void_p(256)      # never dereferenced: the size check comes first
    rc = lib.cugo_b200_chi2_noalloc(0, None, 2, MAX_EDGES + 1, one, one, one, one, 1, cam, 0, C.c_double(1.0),
                                    one, one, one, err, 256)
    assert rc == -1 and b"CUGO_B200_NOALLOC_MAX_EDGES" in err.value
    rc = lib.cugo_b200_chi2_noalloc(0, None, 4, 10, one, one, one, one, 1, cam, 0, C.c_double(1.0),
                                    one, one, one, err, 256)
    assert rc == -1 and b"bad argument" in err.value
