"""Independent numpy statement of the BA model (residuals, SE3 update), used only
to cross-check the C oracle on small graphs by finite differences."""
import numpy as np


def quat_rot(q, X):
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return R @ X


def quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by + ay * bw + az * bx - ax * bz,
                     aw * bz + az * bw + ax * by - ay * bx,
                     aw * bw - ax * bx - ay * by - az * bz])


def se3_exp(d):
    """(omega, upsilon) -> (q, t) of Exp(d), g2o SE3Quat::exp convention."""
    om, up = d[:3], d[3:]
    th = np.linalg.norm(om)
    K = np.array([[0, -om[2], om[1]], [om[2], 0, -om[0]], [-om[1], om[0], 0]])
    if th < 1e-5:
        V = np.eye(3) + 0.5 * K + K @ K / 6
        q = np.array([0.5 * om[0], 0.5 * om[1], 0.5 * om[2], 1.0])
        q /= np.linalg.norm(q)
    else:
        V = np.eye(3) + (1 - np.cos(th)) / th ** 2 * K + (th - np.sin(th)) / th ** 3 * K @ K
        q = np.concatenate([np.sin(th / 2) * om / th, [np.cos(th / 2)]])
    return q, V @ up


def pose_update(pose, d):
    dq, dt = se3_exp(d)
    q = quat_mul(dq, pose[:4])
    q = q / np.linalg.norm(q)
    if q[3] < 0:
        q = -q
    t = dt + quat_rot(dq, pose[4:])
    return np.concatenate([q, t])


def residual(pose, X, meas, cam, dim):
    Xc = quat_rot(pose[:4], X) + pose[4:]
    fx, fy, cx, cy, bf = cam
    u = fx * Xc[0] / Xc[2] + cx
    v = fy * Xc[1] / Xc[2] + cy
    p = np.array([u, v, u - bf / Xc[2]])[:dim]
    return p - meas


def rho(kind, delta, x):
    d2 = delta * delta
    if kind == "none":
        return x, 1.0
    if kind == "huber":
        return (x, 1.0) if x <= d2 else (2 * delta * np.sqrt(x) - d2, delta / np.sqrt(x))
    if kind == "cauchy":
        return d2 * np.log1p(x / d2), 1.0 / (1 + x / d2)
    if kind == "tukey":
        if x <= d2:
            return d2 / 3 * (1 - (1 - x / d2) ** 3), (1 - x / d2) ** 2
        return d2 / 3, 0.0
    raise ValueError(kind)


def dense_system(prob, robust="none", delta=1.0, eps=1e-6):
    """Dense H = J^T W J and b = J^T W e over [poses(6 each) | landmarks(3 each)] of the free
    vertices, J obtained by central differences of the residual under the retraction
    (pose: Exp(d) T, landmark: X + d).  Sign convention of the reference: J = d e / d x,
    b = J^T W e, the step solves H dx = -b ... the reference instead defines J = -d proj
    and solves H dx = b; both give the same dx.  Here: returns (H, b_ref) with b_ref in the
    reference's convention (b_ref = -J^T W e)."""
    nP, nL = prob.nP_act, prob.nL_act
    n = 6 * nP + 3 * nL
    H = np.zeros((n, n))
    b = np.zeros(n)
    chi = 0.0
    for s in prob.sets:
        for i in range(s.n):
            iP, iL = int(s.iP[i]), int(s.iL[i])
            pose, X = prob.poses[iP], prob.landmarks[iL]
            cam = s.cam[i] if s.cam.shape[0] > 1 else s.cam[0]
            w = s.omega[i] if s.omega.shape[0] > 1 else s.omega[0]
            e = residual(pose, X, s.meas[i], cam, s.dim)
            r, dr = rho(robust, delta, w * e @ e)
            chi += r
            W = w * dr
            cols, J = [], []
            if iP < nP:
                Jp = np.zeros((s.dim, 6))
                for k in range(6):
                    d = np.zeros(6); d[k] = eps
                    Jp[:, k] = (residual(pose_update(pose, d), X, s.meas[i], cam, s.dim) -
                                residual(pose_update(pose, -d), X, s.meas[i], cam, s.dim)) / (2 * eps)
                cols += list(range(6 * iP, 6 * iP + 6)); J.append(Jp)
            if iL < nL:
                Jl = np.zeros((s.dim, 3))
                for k in range(3):
                    d = np.zeros(3); d[k] = eps
                    Jl[:, k] = (residual(pose, X + d, s.meas[i], cam, s.dim) -
                                residual(pose, X - d, s.meas[i], cam, s.dim)) / (2 * eps)
                cols += list(range(6 * nP + 3 * iL, 6 * nP + 3 * iL + 3)); J.append(Jl)
            J = np.concatenate(J, axis=1)
            cols = np.array(cols)
            H[np.ix_(cols, cols)] += W * J.T @ J
            b[cols] += -W * J.T @ e
    return H, b, chi
