"""CPU tests (no GPU): the C oracle against closed-form cases, an independent numpy
finite-difference statement of the model, a pure-Python restatement of the reference's host
structure code, and its own invariants."""
import numpy as np
import pytest

import np_ref
import oracle as orc
from cugo_b200 import pack, synth


def tiny_problem(dim=2, robust_outlier=False):
    """2 poses (first fixed at identity) x 3 landmarks, hand-made."""
    poses = np.array([[0, 0, 0, 1, 0, 0, 0], [0, 0.01, 0, 0.99995, -1.0, 0.02, 0.1]], float)
    poses[1, :4] /= np.linalg.norm(poses[1, :4])
    lms = np.array([[0.0, 0.0, 10.0], [1.0, -0.5, 8.0], [-2.0, 0.7, 12.0]])
    cam = np.array([[500.0, 510.0, 320.0, 240.0, 50.0]])
    iP, iL, meas = [], [], []
    for p in range(2):
        for l in range(3):
            e = np_ref.residual(poses[p], lms[l], np.zeros(dim), cam[0], dim)
            iP.append(p); iL.append(l); meas.append(e + 0.3 * np.sin(p + 2 * l + np.arange(dim)))
    # index space: pose 1 is free -> index 0, pose 0 fixed -> index 1
    idx = {1: 0, 0: 1}
    s = pack.EdgeSetArrays(dim=dim, iP=np.array([idx[p] for p in iP], np.int32), iL=np.array(iL, np.int32),
                           meas=np.array(meas), omega=np.array([1.0, 0.5, 2.0, 1.5, 0.7, 1.1]), cam=cam,
                           src=np.arange(6))
    return pack.Problem(poses=poses[[1, 0]].copy(), landmarks=lms, nP_act=1, nL_act=3,
                        pose_ids=np.array([1, 0]), lm_ids=np.arange(3), sets=[s])


def test_projection_known_answer():
    # identity pose, point on the optical axis projects to the principal point
    poses = np.array([[0, 0, 0, 1, 0, 0, 0.0]] * 2)
    lms = np.array([[0.0, 0.0, 5.0]])
    cam = np.array([[500.0, 400.0, 320.0, 240.0, 40.0]])
    s = pack.EdgeSetArrays(dim=3, iP=np.array([0], np.int32), iL=np.array([0], np.int32),
                           meas=np.array([[320.0, 240.0, 320.0 - 40.0 / 5.0]]), omega=np.array([1.0]), cam=cam,
                           src=np.arange(1))
    prob = pack.Problem(poses=poses, landmarks=lms, nP_act=1, nL_act=1, pose_ids=np.arange(2),
                        lm_ids=np.arange(1), sets=[s])
    o = orc.Oracle(prob)
    assert o.compute_errors() == pytest.approx(0.0, abs=1e-20)
    s.meas[0] = [319.0, 242.0, 311.0]
    o = orc.Oracle(prob)
    # e = (1, -2, 1): chi2 = 6
    assert o.compute_errors() == pytest.approx(6.0, rel=1e-14)


@pytest.mark.parametrize("dim", [2, 3])
def test_linear_system_vs_finite_differences(dim):
    prob = tiny_problem(dim)
    o = orc.Oracle(prob)
    o.build_structure()
    chi = o.compute_errors()
    o.build_system()
    H, b, chi_np = np_ref.dense_system(prob)
    assert chi == pytest.approx(chi_np, rel=1e-12)
    nP, nL = 1, 3
    Hpp = o.values("Hpp").reshape(nP, 6, 6).transpose(0, 2, 1)
    Hll = o.values("Hll").reshape(nL, 3, 3).transpose(0, 2, 1)
    Hpl = o.values("Hpl").reshape(-1, 3, 6).transpose(0, 2, 1)       # col-major 6x3
    np.testing.assert_allclose(Hpp[0], H[:6, :6], rtol=1e-6, atol=1e-6 * np.abs(H).max())
    for l in range(nL):
        np.testing.assert_allclose(Hll[l], H[6 + 3 * l:9 + 3 * l, 6 + 3 * l:9 + 3 * l], rtol=1e-6,
                                   atol=1e-6 * np.abs(H).max())
    colptr, rowind = o.structure("hpl_colptr"), o.structure("hpl_rowind")
    for l in range(nL):
        for k in range(colptr[l], colptr[l + 1]):
            assert rowind[k] == 0
            np.testing.assert_allclose(Hpl[k], H[:6, 6 + 3 * l:9 + 3 * l], rtol=1e-6, atol=1e-6 * np.abs(H).max())
    np.testing.assert_allclose(o.values("b"), b, rtol=1e-6, atol=1e-6 * np.abs(b).max())
    # Schur solve == dense damped solve
    lam = 1e-5 * o.max_diagonal()
    assert o.max_diagonal() == pytest.approx(np.diag(H).max(), rel=1e-6)
    o.set_lambda(lam)
    assert o.solve()
    x = np.linalg.solve(H + lam * np.eye(H.shape[0]), b)
    np.testing.assert_allclose(o.values("x"), x, rtol=1e-5, atol=1e-7 * np.abs(x).max())
    # update == numpy retraction
    o.update()
    poses, lms = o.estimates()
    np.testing.assert_allclose(poses[0], np_ref.pose_update(prob.poses[0], o.values("x")[:6]), rtol=0, atol=1e-13)
    np.testing.assert_allclose(poses[1], prob.poses[1])
    np.testing.assert_allclose(lms, prob.landmarks + o.values("x")[6:].reshape(3, 3), atol=1e-14)


@pytest.mark.parametrize("kind,delta", [("huber", 0.8), ("cauchy", 1.0), ("tukey", 3.0)])
def test_robust_kernels_vs_numpy(kind, delta):
    prob = tiny_problem(3)
    o = orc.Oracle(prob, robust=kind, delta=delta)
    o.build_structure()
    chi = o.compute_errors()
    o.build_system()
    H, b, chi_np = np_ref.dense_system(prob, robust=kind, delta=delta)
    assert chi == pytest.approx(chi_np, rel=1e-12)
    np.testing.assert_allclose(o.values("b"), b, rtol=1e-6, atol=1e-6 * np.abs(b).max())


def test_robust_kernel_formulas():
    # reference formulas (cuda_block_solver.cu:985-1027) + g2o Huber, via a 1-edge problem
    poses = np.array([[0, 0, 0, 1, 0, 0, 0.0]] * 2)
    lms = np.array([[0.0, 0.0, 5.0]])
    cam = np.array([[500.0, 400.0, 320.0, 240.0, 40.0]])
    for kind, delta, x, expect in [("huber", 1.0, 0.25, 0.25), ("huber", 1.0, 9.0, 2 * 3 - 1.0),
                                   ("cauchy", 2.0, 4.0, 4 * np.log(2.0)), ("tukey", 2.0, 1.0, 4 / 3 * (1 - 0.75 ** 3)),
                                   ("tukey", 2.0, 9.0, 4 / 3)]:
        s = pack.EdgeSetArrays(dim=2, iP=np.array([0], np.int32), iL=np.array([0], np.int32),
                               meas=np.array([[320.0 - np.sqrt(x), 240.0]]), omega=np.array([1.0]), cam=cam,
                               src=np.arange(1))
        prob = pack.Problem(poses=poses, landmarks=lms, nP_act=1, nL_act=1, pose_ids=np.arange(2),
                            lm_ids=np.arange(1), sets=[s])
        assert orc.Oracle(prob, robust=kind, delta=delta).compute_errors() == pytest.approx(expect, rel=1e-12)


def reference_structure_python(prob):
    """Pure-Python restatement of the reference's host structure code with its dense P x P
    map (src/sparse_block_matrix.cpp:63-262) and Hpl CSC (cuda_block_solver.cu:1571-1604)."""
    nP, nL = prob.nP_act, prob.nL_act
    trip, eid = [], 0
    for s in prob.sets:
        for i in range(s.n):
            if s.iP[i] < nP and s.iL[i] < nL:
                trip.append((int(s.iL[i]), int(s.iP[i]), eid))
            eid += 1
    trip.sort()
    colptr = np.zeros(nL + 1, np.int32)
    for l, _, _ in trip:
        colptr[l + 1] += 1
    colptr = np.cumsum(colptr).astype(np.int32)
    rowind = np.array([p for _, p, _ in trip], np.int32)
    edge2hpl = -np.ones(eid, np.int32)
    for k, (_, _, e) in enumerate(trip):
        edge2hpl[e] = k
    dense = np.zeros((nP, nP), bool)
    nmul = 0
    for l in range(nL):
        idx = sorted(rowind[colptr[l]:colptr[l + 1]])
        for a in range(len(idx)):
            for b2 in range(a, len(idx)):
                dense[idx[a], idx[b2]] = True
                nmul += 1
    rowptr = np.concatenate([[0], np.cumsum(dense.sum(1))]).astype(np.int32)
    colind = np.concatenate([np.flatnonzero(dense[r]) for r in range(nP)]).astype(np.int32)
    # CSR expansion
    rows = [[] for _ in range(6 * nP)]
    for R in range(nP):
        for k in range(rowptr[R], rowptr[R + 1]):
            Cc = colind[k]
            src = 36 * k
            for dc in range(6):
                for dr in range(6):
                    rows[6 * R + dr].append((6 * Cc + dc, src))
                    if R != Cc:
                        rows[6 * Cc + dc].append((6 * R + dr, src))
                    src += 1
    crp = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.int32)
    cci = np.array([c for r in rows for c, _ in r], np.int32)
    b2c = np.array([s2 for r in rows for _, s2 in r], np.int32)
    mul = []
    for l in range(nL):
        for i in range(colptr[l], colptr[l + 1]):
            for j in range(i, colptr[l + 1]):
                R, Cc = rowind[i], rowind[j]
                k = rowptr[R] + int(np.searchsorted(colind[rowptr[R]:rowptr[R + 1]], Cc))
                mul.append((i, j, k))
    mul.sort()
    return dict(hpl_colptr=colptr, hpl_rowind=rowind, edge2hpl=edge2hpl, hsc_rowptr=rowptr, hsc_colind=colind,
                csr_rowptr=crp, csr_colind=cci, bsr2csr=b2c, mulids=np.array(mul, np.int32).reshape(-1))


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("small", {}), ("small", dict(fixed_landmark_frac=0.2, seed=5))])
def test_structure_vs_python_restatement(graphs, name, kw):
    g, prob = graphs(name, **kw)
    o = orc.Oracle(prob)
    o.build_structure()
    ref = reference_structure_python(prob)
    for k, v in ref.items():
        np.testing.assert_array_equal(o.structure(k), v, err_msg=k)
    # CSR columns strictly ascending inside every row
    rp, ci = ref["csr_rowptr"], ref["csr_colind"]
    for r in range(len(rp) - 1):
        assert np.all(np.diff(ci[rp[r]:rp[r + 1]]) > 0)


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_oracle_solve_vs_dense(graphs, name):
    g, prob = graphs(name)
    o = orc.Oracle(prob)
    o.build_structure()
    o.compute_errors()
    o.build_system()
    nP, nL = prob.nP_act, prob.nL_act
    Hpp = o.values("Hpp").reshape(nP, 6, 6).transpose(0, 2, 1)
    Hll = o.values("Hll").reshape(nL, 3, 3).transpose(0, 2, 1)
    Hpl = o.values("Hpl").reshape(-1, 3, 6).transpose(0, 2, 1)
    n = 6 * nP + 3 * nL
    H = np.zeros((n, n))
    for p in range(nP):
        H[6 * p:6 * p + 6, 6 * p:6 * p + 6] = Hpp[p]
    for l in range(nL):
        H[6 * nP + 3 * l:6 * nP + 3 * l + 3, 6 * nP + 3 * l:6 * nP + 3 * l + 3] = Hll[l]
    cp, ri = o.structure("hpl_colptr"), o.structure("hpl_rowind")
    for l in range(nL):
        for k in range(cp[l], cp[l + 1]):
            p = ri[k]
            H[6 * p:6 * p + 6, 6 * nP + 3 * l:6 * nP + 3 * l + 3] = Hpl[k]
            H[6 * nP + 3 * l:6 * nP + 3 * l + 3, 6 * p:6 * p + 6] = Hpl[k].T
    b = o.values("b")
    lam = 1e-5 * o.max_diagonal()
    o.set_lambda(lam)
    assert o.solve()
    x = np.linalg.solve(H + lam * np.eye(n), b)
    np.testing.assert_allclose(o.values("x"), x, rtol=1e-7, atol=1e-9 * np.abs(x).max())
    assert o.compute_scale(lam) == pytest.approx(float(x @ (lam * x + b)), rel=1e-7)


def test_lm_decreases_and_matches_truth(graphs):
    g, prob = graphs("small")
    o = orc.Oracle(prob)
    chi2, lam, trials = o.optimize(10)
    assert len(chi2) == 10 and np.all(np.diff(chi2) < 0)
    assert np.all(lam > 0)


def test_lm_control_flow_rejects(graphs):
    # a start far from the optimum provokes rejected trials; statistics stay consistent
    g, prob = graphs("small", seed=11)
    prob2 = pack.Problem(poses=prob.poses.copy(), landmarks=prob.landmarks + 3.0, nP_act=prob.nP_act,
                         nL_act=prob.nL_act, pose_ids=prob.pose_ids, lm_ids=prob.lm_ids, sets=prob.sets)
    o = orc.Oracle(prob2)
    F0 = o.compute_errors()
    chi2, lam, trials = o.optimize(15)
    assert chi2[0] <= F0
    assert np.all(np.diff(chi2) <= 0)


def test_float32_oracle_close_to_double(graphs):
    g, prob = graphs("small")
    c64, _, _ = orc.Oracle(prob).optimize(5)
    c32, _, _ = orc.Oracle(prob, f32=True).optimize(5)
    n = min(len(c64), len(c32))
    np.testing.assert_allclose(c32[:n], c64[:n], rtol=1e-3)
