"""depth edges, pose-only ICP plane / line edges, graphs without free landmarks.

The reference's own versions of these are not a model to pin to (depth: residual sign against its Jacobian,
src/cuda/cuda_block_solver.cu:2341-2343; line: Jacobian through an uninitialised matrix, :2286-2288; line chi2 without the
square, :2407-2409), so PARITY OF THE JACOBIANS / LM STEPS IS UNPINNED against the reference for these edge kinds.  Checked:
  * the residuals against the reference itself (goldens ref_f4_*, bottom of this file): per edge, chi2, CPU oracle and GPU;
  * CPU: the oracle's right-hand side b = J^T w e against central finite differences of its own chi2 for every kind
    (residuals are the reference's: camProjectDepth :426-432, signedDistance :372-377, distance :352-370), and that LM on
    these graphs reaches the noise floor;
  * GPU: the CUDA engine through the C ABI against the oracle -- stage values, chi2 per iteration within 1e-9 relative,
    identical trial counts, estimates.
"""
import copy

import numpy as np
import pytest

CHI2_RTOL = 1e-9
VALUE_RTOL = 1e-10


def relerr(a, b):
    s = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / s


def _with_pose_step(prob, p, k, h):
    """prob with pose p <- Exp(h e_k) pose p (the engine's update, reference updatePosesKernel :1444-1455), to first
    order in the translation part -- enough for a central difference."""
    from cugo_b200 import synth
    out = copy.copy(prob)
    poses = prob.poses.copy()
    d = np.zeros(6); d[k] = h
    R = synth.quat_to_rot(poses[p, :4]); t = poses[p, 4:]
    dR = synth._rodrigues(d[:3][None])[0]
    K = np.array([[0, -d[2], d[1]], [d[2], 0, -d[0]], [-d[1], d[0], 0]])
    poses[p, :4] = synth.rot_to_quat((dR @ R)[None])[0]
    poses[p, 4:] = dR @ t + (np.eye(3) + 0.5 * K) @ d[3:]
    out.poses = poses
    return out


def _with_lm_step(prob, l, k, h):
    out = copy.copy(prob)
    lms = prob.landmarks.copy()
    lms[l, k] += h
    out.landmarks = lms
    return out


def _problems():
    from cugo_b200 import pack, synth
    ba = pack.pack(synth.make_graph("tiny"))
    return {
        "plane": synth.make_icp_problem(n_lines=0, rot=0.2, trans=0.5),
        "line": synth.make_icp_problem(n_planes=0, rot=0.2, trans=0.5),
        "depth": synth.stereo_to_depth(ba),
        "depth_pose_only": synth.fix_all_landmarks(synth.stereo_to_depth(ba)),
    }


@pytest.mark.parametrize("kind", ["plane", "line", "depth", "depth_pose_only"])
@pytest.mark.parametrize("robust", ["none", "cauchy"])
def test_oracle_rhs_is_the_chi2_gradient(kind, robust):
    """b = J^T w e with J = -de/dx and w = omega rho'(omega e^2)  <=>  b = -1/2 d(chi2)/dx."""
    import oracle
    prob = _problems()[kind]
    kw = dict(robust=robust, delta=0.7)
    o = oracle.Oracle(prob, **kw)
    o.build_structure(); o.compute_errors(); o.build_system()
    b = o.values("b")
    h = 1e-6
    for p in range(min(prob.nP_act, 3)):
        for k in range(6):
            g = (oracle.Oracle(_with_pose_step(prob, p, k, h), **kw).compute_errors()
                 - oracle.Oracle(_with_pose_step(prob, p, k, -h), **kw).compute_errors()) / (2 * h)
            assert abs(b[p * 6 + k] + 0.5 * g) < 1e-6 * max(1.0, abs(g)), (kind, p, k)
    for l in range(min(prob.nL_act, 3)):
        for k in range(3):
            g = (oracle.Oracle(_with_lm_step(prob, l, k, h), **kw).compute_errors()
                 - oracle.Oracle(_with_lm_step(prob, l, k, -h), **kw).compute_errors()) / (2 * h)
            assert abs(b[prob.nP_act * 6 + l * 3 + k] + 0.5 * g) < 1e-6 * max(1.0, abs(g)), (kind, "lm", l, k)


def test_oracle_icp_reaches_the_noise_floor():
    """Noise-free matches: LM brings chi2 to (numerically) zero from a 0.2 rad / 0.5 m start -- the Jacobians point the
    right way (the reference's depth edge would climb instead)."""
    import oracle
    from cugo_b200 import synth
    prob = synth.make_icp_problem(n_poses=4, n_planes=300, n_lines=0, noise=0.0, rot=0.2, trans=0.5)
    o = oracle.Oracle(prob)
    o.build_structure()
    f0 = o.compute_errors()
    chi, _, _ = o.optimize(20)
    assert f0 > 1.0 and chi[-1] < 1e-12 * f0


def test_oracle_pose_only_ba_matches_full_ba_with_frozen_landmarks():
    """A graph whose landmarks are all fixed takes the block-diagonal solve; one Gauss-Newton step from the same
    estimates equals the pose part of ... itself computed through the dense normal equations in numpy."""
    import oracle
    from cugo_b200 import pack, synth
    prob = synth.fix_all_landmarks(pack.pack(synth.make_graph("tiny")))
    o = oracle.Oracle(prob)
    o.build_structure(); o.compute_errors(); o.build_system()
    H = o.values("Hpp").reshape(-1, 6, 6)
    b = o.values("b").reshape(-1, 6)
    lam = 1e-5 * o.max_diagonal()
    o.set_lambda(lam)
    assert o.solve()
    x = o.values("x").reshape(-1, 6)
    for p in range(prob.nP_act):
        np.testing.assert_allclose(x[p], np.linalg.solve(H[p].T + lam * np.eye(6), b[p]), rtol=1e-9, atol=1e-14)


# ---------------------------------------------------------------------------------------------------------------------
# GPU: the CUDA engine (C ABI) against the oracle
# ---------------------------------------------------------------------------------------------------------------------
def _pair(prob, **kw):
    import oracle
    from cugo_b200 import solver
    s = solver.BlockSolver(0)
    s.initialize(prob, **kw)
    return s, oracle.Oracle(prob, **kw)


def _check_trajectory(s, o, niter, est_tol=1e-8):
    chi_g, lam_g, tr_g = s.optimize(niter)
    chi_o, lam_o, tr_o = o.optimize(niter)
    assert len(chi_g) == len(chi_o)
    np.testing.assert_array_equal(tr_g, tr_o)
    np.testing.assert_allclose(chi_g, chi_o, rtol=CHI2_RTOL, atol=1e-18)
    np.testing.assert_allclose(lam_g, lam_o, rtol=1e-6)
    pg, lg = s.estimates(); po, lo = o.estimates()
    assert np.sqrt(np.mean((pg - po) ** 2)) < est_tol
    if lg.size:
        assert np.sqrt(np.mean((lg - lo) ** 2)) < est_tol
    return chi_g, tr_g


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", [("tiny", {}), ("small", {}), ("small", dict(per_edge_camera=True)),
                                     ("small", dict(fixed_landmark_frac=0.15, seed=21))])
def test_depth_edges_stages_and_trajectory(graphs, name, kw):
    """DepthEdgeSet beside a mono set in a graph with free landmarks: every stage of the Schur path, then LM."""
    from cugo_b200 import synth
    prob = synth.stereo_to_depth(graphs(name, **kw)[1])
    assert any(getattr(st, "kind", "") == "depth" for st in prob.sets)
    s, o = _pair(prob)
    s.build_structure(); o.build_structure()
    assert s.compute_errors() == pytest.approx(o.compute_errors(), rel=1e-12)
    s.build_system(); o.build_system()
    for k in ("Hpp", "Hll", "Hpl", "b"):
        assert relerr(s.values(k), o.values(k)) < VALUE_RTOL, k
    lam = 1e-5 * o.max_diagonal()
    assert s.max_diagonal() == pytest.approx(o.max_diagonal(), rel=1e-12)
    s.set_lambda(lam); o.set_lambda(lam)
    assert s.solve() and o.solve()
    assert relerr(s.values("x"), o.values("x")) < 1e-8
    s2, o2 = _pair(prob, robust="huber", delta=2.0)
    _check_trajectory(s2, o2, 8)


@pytest.mark.gpu
@pytest.mark.parametrize("robust", ["none", "cauchy"])
@pytest.mark.parametrize("depth", [False, True])
def test_graph_without_free_landmarks(graphs, robust, depth):
    """All landmarks fixed (the reference's doSchur_ == false path): block-diagonal solve, stage by stage and as LM."""
    from cugo_b200 import synth
    prob = graphs("small")[1]
    if depth:
        prob = synth.stereo_to_depth(prob)
    prob = synth.fix_all_landmarks(prob)
    kw = dict(robust=robust, delta=3.0)
    s, o = _pair(prob, **kw)
    s.build_structure(); o.build_structure()
    assert s.compute_errors() == pytest.approx(o.compute_errors(), rel=1e-12)
    s.build_system(); o.build_system()
    for k in ("Hpp", "b"):
        assert relerr(s.values(k), o.values(k)) < VALUE_RTOL, k
    assert s.values("Hll").size == 0 and s.values("Hpl").size == 0 and s.values("Hsc").size == 0
    lam = 1e-5 * o.max_diagonal()
    assert s.max_diagonal() == pytest.approx(o.max_diagonal(), rel=1e-12)
    s.set_lambda(lam); o.set_lambda(lam)
    assert s.solve() and o.solve()
    assert relerr(s.values("x"), o.values("x")) < 1e-9
    assert s.compute_scale(lam) == pytest.approx(o.compute_scale(lam), rel=1e-9)
    s.update(); o.update()
    assert np.abs(s.estimates()[0] - o.estimates()[0]).max() < 1e-10
    assert s.compute_errors() == pytest.approx(o.compute_errors(), rel=1e-9)
    s2, o2 = _pair(prob, **kw)
    _check_trajectory(s2, o2, 10)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["plane", "line", "both", "one_pose_many_edges", "with_fixed_pose", "cauchy", "far_start"])
def test_icp_edge_sets(case):
    """PlaneEdgeSet / LineEdgeSet: pose-only graphs of scalar residuals."""
    from cugo_b200 import synth
    kw = {}
    if case == "plane":
        prob = synth.make_icp_problem(n_lines=0, rot=0.2, trans=0.5)
    elif case == "line":
        prob = synth.make_icp_problem(n_planes=0, rot=0.2, trans=0.5)
    elif case == "both":
        prob = synth.make_icp_problem(n_poses=40, n_planes=6000, n_lines=5000, rot=0.1, trans=0.3)
    elif case == "one_pose_many_edges":      # one scan against a map: several slices of one pose
        prob = synth.make_icp_problem(n_poses=1, n_planes=9000, n_lines=7000, rot=0.1, trans=0.3)
    elif case == "with_fixed_pose":
        prob = synth.make_icp_problem(n_poses=5, n_fixed=2, rot=0.1, trans=0.3)
    elif case == "cauchy":
        prob = synth.make_icp_problem(rot=0.3, trans=1.0)
        kw = dict(robust="cauchy", delta=0.5)
    else:                                    # LM has to reject trials
        prob = synth.make_icp_problem(n_poses=3, n_planes=60, n_lines=60, rot=1.2, trans=6.0, seed=5)
    s, o = _pair(prob, **kw)
    s.build_structure(); o.build_structure()
    assert s.compute_errors() == pytest.approx(o.compute_errors(), rel=1e-12)
    s.build_system(); o.build_system()
    for k in ("Hpp", "b"):
        assert relerr(s.values(k), o.values(k)) < VALUE_RTOL, k
    assert s.max_diagonal() == pytest.approx(o.max_diagonal(), rel=1e-12)
    s2, o2 = _pair(prob, **kw)
    chi, tr = _check_trajectory(s2, o2, 15)
    if case == "far_start":
        assert tr.sum() >= 1


@pytest.mark.gpu
def test_icp_edges_beside_a_schur_graph(graphs):
    """ICP sets only add to Hpp / bp of their poses, so they may sit beside BA sets in a graph with free landmarks."""
    import dataclasses
    from cugo_b200 import synth
    ba = graphs("small")[1]
    icp = synth.make_icp_problem(n_poses=ba.nP_act, n_planes=500, n_lines=400, rot=0.0, trans=0.0, noise=0.05,
                                 true_poses=ba.poses)
    prob = dataclasses.replace(ba, sets=[ba.sets[0], icp.sets[0]] + list(ba.sets[1:]) + [icp.sets[1]])
    assert len(prob.sets) == 4
    s, o = _pair(prob)
    s.build_structure(); o.build_structure()
    assert s.compute_errors() == pytest.approx(o.compute_errors(), rel=1e-12)
    s.build_system(); o.build_system()
    for k in ("Hpp", "Hll", "Hpl", "b"):
        assert relerr(s.values(k), o.values(k)) < VALUE_RTOL, k
    assert s.max_diagonal() == pytest.approx(o.max_diagonal(), rel=1e-12)
    s2, o2 = _pair(prob, robust=["none", "huber", "none", "cauchy"], delta=0.2)
    _check_trajectory(s2, o2, 8)


@pytest.mark.gpu
def test_icp_argument_errors():
    from cugo_b200 import solver, synth
    prob = synth.make_icp_problem(n_poses=3, n_fixed=1)
    bad = copy.copy(prob)
    bad.sets = [copy.copy(prob.sets[0])]
    bad.sets[0].iP = np.full_like(prob.sets[0].iP, 2)          # the fixed pose
    s = solver.BlockSolver(0)
    with pytest.raises(solver.CugoError):
        s.initialize(bad)
    s.initialize(prob)
    with pytest.raises(solver.CugoError):                        # the reference's ICP kernels know no outlier mask
        s.set_outlier_threshold(0, 1.0)
    with pytest.raises(solver.CugoError):                        # no Schur structure to read
        s.structure("hsc_rowptr")
    s.optimize(2)


@pytest.mark.gpu
@pytest.mark.parametrize("rk,delta", [(0, 1.0), (1, 0.5)])
def test_cpp_facade_icp_sample_matches_binding(tmp_path, rk, delta):
    """PlaneEdgeSet / LineEdgeSet / PoseVertexSet through include/cugo_b200.hpp (samples/sample_icp.cpp) drive the same
    engine as the ctypes binding: identical chi2 per iteration and final poses."""
    import json
    import os
    import subprocess
    from cugo_b200 import graphio, solver, synth
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "cuda-bundle-adjustment_b200", "lib", "sample_icp")
    prob = synth.make_icp_problem(n_poses=7, n_fixed=2, n_planes=900, n_lines=700, rot=0.2, trans=0.6, seed=3)
    path, est = str(tmp_path / "p.icp"), str(tmp_path / "poses.bin")
    graphio.write_icp(prob, path)
    r = subprocess.run([exe, path, "8", str(rk), repr(delta), est], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    s = solver.BlockSolver(0)
    s.initialize(prob, robust={0: "none", 1: "cauchy"}[rk], delta=delta)
    chi, _, _ = s.optimize(8)
    assert out["chi2"] == [float(c) for c in chi]
    np.testing.assert_array_equal(np.fromfile(est, np.float64).reshape(-1, 7), s.estimates()[0])


@pytest.mark.gpu
def test_outlier_gating_without_free_landmarks(graphs):
    """Row f3 on the block-diagonal path: threshold -> mask on the device at the end of optimize(), same masks and chi2 as
    the oracle's restatement of the reference flow."""
    from cugo_b200 import synth
    prob = synth.fix_all_landmarks(graphs("small", outlier_frac=0.05, seed=7002)[1])
    s, o = _pair(prob)
    for e in (s, o):
        for i in range(len(prob.sets)):
            e.set_outlier_threshold(i, 30.0)
    chi_g, _, tr_g = s.optimize(4); chi_o, _, tr_o = o.optimize(4)
    np.testing.assert_array_equal(tr_g, tr_o)
    np.testing.assert_allclose(chi_g, chi_o, rtol=CHI2_RTOL)
    total = 0
    for i in range(len(prob.sets)):
        mg, cg = s.outliers(i); mo, co = o.outliers(i)
        np.testing.assert_array_equal(mg, mo)
        assert cg == co
        total += cg
    assert total > 0
    chi_g2, _, tr_g2 = s.optimize(4); chi_o2, _, tr_o2 = o.optimize(4)
    np.testing.assert_array_equal(tr_g2, tr_o2)
    np.testing.assert_allclose(chi_g2, chi_o2, rtol=CHI2_RTOL)
    assert chi_g2[0] < chi_g[-1]


# ---------------------------------------------------------------------------------------------------------------------
# What the reference CAN pin for these edge kinds: its residuals (tests/golden/ref_f4_*: the unmodified reference run on
# B200 by oracle/run_ref_goldens_f4.py through oracle/ref_driver_f4.cpp)
# ---------------------------------------------------------------------------------------------------------------------
def _golden(name):
    import json
    import os
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    return json.load(open(os.path.join(d, name + ".json"))), np.load(os.path.join(d, name + ".npz"))


def _depth_small():
    from cugo_b200 import pack, synth
    return synth.stereo_to_depth(pack.pack(synth.make_graph("small")))


def test_depth_residuals_match_reference_golden():
    """DepthEdgeSet: chi2 of computeErrors() equals the reference's to 1e-12 and every residual is MINUS the reference's
    (it takes meas - proj, cuda_block_solver.cu:2341-2343; ours proj - meas like the other edges).  The golden also records
    what the reference's LM does with that sign and the stereo Jacobian: all trials of the first iteration rejected
    ("levenberg iterations = 9", rho < 0, lambda blown up to 2e19), chi2 never moves."""
    import oracle
    meta, z = _golden("ref_f4_depth_small")
    prob = _depth_small()
    o = oracle.Oracle(prob)
    o.build_structure()
    assert o.compute_errors() == pytest.approx(meta["F0"], rel=1e-12)
    si = [i for i, s in enumerate(prob.sets) if s.kind == "depth"][0]
    s = prob.sets[si]
    err, _ = o.edge_errors(si)
    at = {(int(a), int(b)): k for k, (a, b) in enumerate(zip(s.iP, s.iL))}
    idx = np.array([at[(int(a), int(b))] for a, b in z["edge2pl"]])
    assert len(idx) == s.n
    np.testing.assert_allclose(err[idx], -z["errors"], rtol=0, atol=1e-10)
    assert meta["chi2"] == [meta["F0"]] and "levenberg iterations = 9" in meta["trials_log"][0]
    chi, _, tr = oracle.Oracle(prob).optimize(5)         # with the derived Jacobian the same graph converges
    assert len(chi) == 5 and tr.sum() == 0 and chi[-1] < 0.03 * meta["F0"]


def test_icp_residuals_match_reference_golden():
    """PlaneEdgeSet / LineEdgeSet: per-edge residuals equal the reference's (1e-13); its plane chi2 is sum omega e^2 as
    ours, its line "chi2" is the plain sum of e (cuda_block_solver.cu:2407-2409) -- pinned as such."""
    import oracle
    from cugo_b200 import synth
    meta, z = _golden("ref_f4_icp")
    prob = synth.make_icp_problem(**meta["icp_kw"])
    o = oracle.Oracle(prob)
    o.build_structure(); o.compute_errors()
    for si, k in ((0, "plane"), (1, "line")):
        s = prob.sets[si]
        assert s.kind == k
        err, chi = o.edge_errors(si)
        at = {s.meas[i].tobytes(): i for i in range(s.n)}
        idx = np.array([at[z[k + "_meas"][i].tobytes()] for i in range(z[k + "_meas"].shape[0])])
        assert len(idx) == s.n and np.array_equal(s.iP[idx], z[k + "_pose"])
        np.testing.assert_allclose(err[idx, 0], z[k + "_errors"], rtol=0, atol=1e-13)
        om = s.omega if s.omega.shape[0] == s.n else np.full(s.n, s.omega[0])
        if k == "plane":
            assert chi.sum() == pytest.approx(meta["F_plane"], rel=1e-13)
            assert (om * err[:, 0] ** 2).sum() == pytest.approx(meta["F_plane"], rel=1e-13)
        else:
            assert err[:, 0].sum() == pytest.approx(meta["F_line"], rel=1e-13)


@pytest.mark.gpu
def test_engine_chi2_matches_reference_goldens_f4():
    """The CUDA engine's computeErrors against the reference's numbers directly: depth graph, plane set alone, and the line
    set alone against sum omega e_ref^2 built from the reference's own per-edge residuals."""
    import dataclasses
    from cugo_b200 import solver, synth
    meta, _ = _golden("ref_f4_depth_small")
    s = solver.BlockSolver(0)
    s.initialize(_depth_small())
    assert s.compute_errors() == pytest.approx(meta["F0"], rel=1e-12)
    meta, z = _golden("ref_f4_icp")
    prob = synth.make_icp_problem(**meta["icp_kw"])
    s = solver.BlockSolver(0)
    s.initialize(dataclasses.replace(prob, sets=[prob.sets[0]]))
    assert s.compute_errors() == pytest.approx(meta["F_plane"], rel=1e-12)
    ln = prob.sets[1]
    at = {ln.meas[i].tobytes(): i for i in range(ln.n)}
    idx = np.array([at[z["line_meas"][i].tobytes()] for i in range(ln.n)])
    om = ln.omega if ln.omega.shape[0] == ln.n else np.full(ln.n, ln.omega[0])
    s = solver.BlockSolver(0)
    s.initialize(dataclasses.replace(prob, sets=[ln]))
    assert s.compute_errors() == pytest.approx(float((om[idx] * z["line_errors"] ** 2).sum()), rel=1e-12)
