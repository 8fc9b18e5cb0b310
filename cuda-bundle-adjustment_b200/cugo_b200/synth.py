"""Seeded synthetic bundle-adjustment graphs.

The reference's sample inputs (samples/ba_input.7z) are absent from the snapshot
(/root/reference/.MISSING_LARGE_BLOBS:1), so every workload is generated.  The
record layout mirrors the reference JSON schema read by
samples/sample_ba_from_file/main.cpp:88-160 (pose_vertices{id,fixed,q,t},
landmark_vertices{id,fixed,Xw}, monocular_edges / stereo_edges
{vertexP,vertexL,measurement,information}, fx fy cx cy bf) but is held as flat
numpy arrays keyed by vertex *id*.

Geometry: a vehicle drives laps of a circuit (period `loop_period` frames, 1 m
spacing) looking along the tangent; pose (q,t) maps world -> camera,
Xc = R(q) Xw + t (reference cuda_block_solver.cu:396-402).  Every landmark is
seen by `d` consecutive frames from its anchor frame and, with probability
`loop_frac`, by a few frames one lap later (loop closure => off-band Hsc blocks).
"""
from __future__ import annotations

import dataclasses
import numpy as np

CAMERA = dict(fx=718.856, fy=718.856, cx=607.1928, cy=185.2157, bf=386.1448)

# name -> (poses, landmarks, edges, stereo edges, seed, loop_period)
SHAPES = {
    "tiny": (12, 90, 400, 240, 7100, 0),
    "small": (40, 600, 2600, 1600, 7101, 0),
    "loop": (300, 4000, 18000, 6000, 7102, 150),
    "kitti07": (248, 26127, 95037, 0, 7001, 0),
    "kitti00": (1332, 133383, 561116, 336672, 7000, 900),
    "venice": (1778, 993923, 5001946, 0, 7003, 0),
    "city": (20000, 5000000, 30000000, 0, 7004, 4000),      # BASELINE.json configs[4]
}


@dataclasses.dataclass
class Graph:
    """Flat, id-keyed graph.  Arrays are float64 / int32 / uint8."""
    pose_q: np.ndarray      # [P,4] (x,y,z,w)
    pose_t: np.ndarray      # [P,3]
    pose_fixed: np.ndarray  # [P] uint8
    lm_xyz: np.ndarray      # [L,3]
    lm_fixed: np.ndarray    # [L] uint8
    mono_p: np.ndarray      # [Em] pose id
    mono_l: np.ndarray      # [Em] landmark id
    mono_meas: np.ndarray   # [Em,2]
    mono_info: np.ndarray   # [Em]
    stereo_p: np.ndarray
    stereo_l: np.ndarray
    stereo_meas: np.ndarray  # [Es,3]
    stereo_info: np.ndarray
    camera: np.ndarray      # [5] fx fy cx cy bf
    name: str = ""
    # ground truth (not part of the reference schema; used by tests only)
    gt_pose_q: np.ndarray | None = None
    gt_pose_t: np.ndarray | None = None
    gt_lm_xyz: np.ndarray | None = None

    @property
    def n_poses(self):
        return self.pose_q.shape[0]

    @property
    def n_landmarks(self):
        return self.lm_xyz.shape[0]

    @property
    def n_edges(self):
        return self.mono_p.shape[0] + self.stereo_p.shape[0]


def quat_to_rot(q):
    """(x,y,z,w) -> R, batched.  Same convention as reference
    cuda_block_solver.cu:449-478."""
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z)
    R[..., 0, 1] = 2 * (x * y - z * w)
    R[..., 0, 2] = 2 * (x * z + y * w)
    R[..., 1, 0] = 2 * (x * y + z * w)
    R[..., 1, 1] = 1 - 2 * (x * x + z * z)
    R[..., 1, 2] = 2 * (y * z - x * w)
    R[..., 2, 0] = 2 * (x * z - y * w)
    R[..., 2, 1] = 2 * (y * z + x * w)
    R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def rot_to_quat(R):
    """Batched R -> (x,y,z,w) with w >= 0 (numerically safe branch per row)."""
    R = np.asarray(R)
    out = np.empty(R.shape[:-2] + (4,))
    flat_R = R.reshape(-1, 3, 3)
    flat_o = out.reshape(-1, 4)
    for n in range(flat_R.shape[0]):
        m = flat_R[n]
        tr = m[0, 0] + m[1, 1] + m[2, 2]
        if tr > 0:
            s = np.sqrt(tr + 1.0) * 2
            w = 0.25 * s
            x = (m[2, 1] - m[1, 2]) / s
            y = (m[0, 2] - m[2, 0]) / s
            z = (m[1, 0] - m[0, 1]) / s
        else:
            i = int(np.argmax([m[0, 0], m[1, 1], m[2, 2]]))
            j, k = (i + 1) % 3, (i + 2) % 3
            s = np.sqrt(m[i, i] - m[j, j] - m[k, k] + 1.0) * 2
            v = [0.0, 0.0, 0.0]
            v[i] = 0.25 * s
            w = (m[k, j] - m[j, k]) / s
            v[j] = (m[j, i] + m[i, j]) / s
            v[k] = (m[k, i] + m[i, k]) / s
            x, y, z = v
        q = np.array([x, y, z, w])
        if w < 0:
            q = -q
        flat_o[n] = q / np.linalg.norm(q)
    return out


def _rodrigues(rv):
    """Batched rotation vector -> rotation matrix."""
    th = np.linalg.norm(rv, axis=-1)
    K = np.zeros(rv.shape[:-1] + (3, 3))
    K[..., 0, 1], K[..., 0, 2] = -rv[..., 2], rv[..., 1]
    K[..., 1, 0], K[..., 1, 2] = rv[..., 2], -rv[..., 0]
    K[..., 2, 0], K[..., 2, 1] = -rv[..., 1], rv[..., 0]
    th2 = np.maximum(th, 1e-12)
    a = (np.sin(th2) / th2)[..., None, None]
    b = ((1 - np.cos(th2)) / (th2 * th2))[..., None, None]
    return np.eye(3) + a * K + b * (K @ K)


def _degrees(rng, L, E, dmin=2, dmax=14):
    """Per-landmark track lengths with sum exactly E."""
    mean_extra = E / L - dmin
    if mean_extra < 0:
        raise ValueError("E/L must be >= %d" % dmin)
    d = dmin + rng.geometric(1.0 / (1.0 + mean_extra), size=L) - 1
    d = np.clip(d, dmin, dmax)
    diff = int(E - d.sum())
    # fix up the total by +-1 steps on random landmarks that have room
    while diff != 0:
        step = 1 if diff > 0 else -1
        ok = np.flatnonzero((d < dmax) if step > 0 else (d > dmin))
        take = min(abs(diff), ok.size)
        if take == 0:
            raise ValueError("cannot reach requested edge count")
        pick = rng.choice(ok, size=take, replace=False)
        d[pick] += step
        diff -= step * take
    return d.astype(np.int64)


def make_graph(name="small", *, n_poses=None, n_landmarks=None, n_edges=None,
               n_stereo=None, seed=None, loop_period=None, loop_frac=0.03,
               outlier_frac=0.0, fix_first=2, noise_px=1.0,
               fixed_landmark_frac=0.0) -> Graph:
    P0, L0, E0, S0, seed0, loop0 = SHAPES.get(name, SHAPES["small"])
    P = n_poses or P0
    L = n_landmarks or L0
    E = n_edges or E0
    Es = S0 if n_stereo is None else n_stereo
    seed = seed0 if seed is None else seed
    loop_period = loop0 if loop_period is None else loop_period
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, bf = (CAMERA[k] for k in ("fx", "fy", "cx", "cy", "bf"))

    # ---- ground-truth trajectory: laps of a circuit, 1 m per frame ----------
    period = loop_period if loop_period > 0 else max(4 * P, 400)
    radius = period / (2 * np.pi)
    f = np.arange(P)
    ang = 2 * np.pi * f / period
    lap = f // period
    r = radius + 0.35 * lap                      # laps do not coincide exactly
    cam_pos = np.stack([r * np.cos(ang), 0.02 * np.sin(0.05 * f), r * np.sin(ang)], -1)
    fwd = np.stack([-np.sin(ang), np.zeros(P), np.cos(ang)], -1)   # tangent
    up = np.tile(np.array([0.0, -1.0, 0.0]), (P, 1))               # camera y points down
    right = np.cross(up, fwd)
    # rows of R (world->camera) are the camera axes expressed in the world
    Rwc = np.stack([right, up, fwd], axis=1)
    twc = -(Rwc @ cam_pos[..., None])[..., 0]

    # ---- tracks ------------------------------------------------------------
    d = _degrees(rng, L, E, dmax=min(14, P))
    n_loop_obs = np.zeros(L, np.int64)
    if loop_period > 0:
        cand = rng.random(L) < loop_frac
        n_loop_obs = np.where(cand, np.minimum(d // 2, 3), 0)
    d_local = d - n_loop_obs
    anchor = np.empty(L, np.int64)
    hi = np.maximum(P - d_local, 0)
    anchor[:] = np.floor(rng.random(L) * (hi + 1)).astype(np.int64)
    anchor = np.minimum(anchor, hi)
    # a landmark with loop observations needs frames one lap later to exist
    if loop_period > 0:
        too_late = (n_loop_obs > 0) & (anchor + loop_period + n_loop_obs > P)
        lo_hi = np.maximum(P - loop_period - n_loop_obs, 0)
        anchor = np.where(too_late, np.minimum(anchor % max(loop_period, 1), lo_hi), anchor)
    order = np.argsort(anchor, kind="stable")      # ids grow with time, as in SLAM
    anchor, d, d_local, n_loop_obs = anchor[order], d[order], d_local[order], n_loop_obs[order]

    # edge list (pose id, landmark id) in landmark-major order
    lm_of_edge = np.repeat(np.arange(L), d)
    start = np.cumsum(d) - d
    k_in_track = np.arange(E) - np.repeat(start, d)
    dl_rep = np.repeat(d_local, d)
    pose_of_edge = np.repeat(anchor, d) + np.where(
        k_in_track < dl_rep, k_in_track, loop_period + (k_in_track - dl_rep))
    assert pose_of_edge.max() < P and pose_of_edge.min() >= 0

    # landmark position in its anchor camera frame; landmarks that would end up
    # with Xc.z <= 1 in some observing frame (tight turns) are re-drawn closer
    # to the optical axis until every observation is in front of its camera
    Ra, ta = Rwc[anchor], twc[anchor]
    lm_gt = np.empty((L, 3))
    todo = np.arange(L)
    lat = 0.45
    for _ in range(12):
        n = todo.size
        depth = rng.uniform(np.maximum(5.0, d_local[todo] + 4.0), 60.0)
        u = (0.5 + rng.uniform(-lat, lat, n)) * 2 * cx
        v = (0.5 + rng.uniform(-lat, lat, n)) * 2 * cy
        Xc = np.stack([(u - cx) / fx * depth, (v - cy) / fy * depth, depth], -1)
        lm_gt[todo] = (np.transpose(Ra[todo], (0, 2, 1)) @ (Xc - ta[todo])[..., None])[..., 0]
        z_e = ((Rwc[pose_of_edge] @ lm_gt[lm_of_edge][..., None])[..., 0] + twc[pose_of_edge])[:, 2]
        zmin = np.full(L, np.inf)
        np.minimum.at(zmin, lm_of_edge, z_e)
        todo = np.flatnonzero(zmin <= 1.5)
        if todo.size == 0:
            break
        lat *= 0.6

    # ---- measurements -------------------------------------------------------
    Xc_e = (Rwc[pose_of_edge] @ lm_gt[lm_of_edge][..., None])[..., 0] + twc[pose_of_edge]
    z = Xc_e[:, 2]
    if not np.all(z > 1.0):
        # landmarks that would fall behind a loop-closing camera: move them to a
        # safe depth band relative to the *worst* observing frame
        raise AssertionError("generator produced a point behind a camera (z<=1)")
    pu = fx * Xc_e[:, 0] / z + cx
    pv = fy * Xc_e[:, 1] / z + cy
    pr = pu - bf / z
    noise = rng.normal(0.0, noise_px, (E, 3))
    meas = np.stack([pu, pv, pr], -1) + noise
    if outlier_frac > 0:
        bad = rng.random(E) < outlier_frac
        mag = rng.uniform(20.0, 100.0, E) * rng.choice([-1.0, 1.0], E)
        meas[bad, 0] += mag[bad]
        meas[bad, 1] -= 0.5 * mag[bad]
    info = 1.2 ** (-2.0 * rng.integers(0, 8, E))

    is_stereo = np.zeros(E, bool)
    is_stereo[rng.permutation(E)[:Es]] = True
    shuffle = rng.permutation(E)                  # edge arrival order is arbitrary
    pose_of_edge, lm_of_edge, meas, info, is_stereo = (
        a[shuffle] for a in (pose_of_edge, lm_of_edge, meas, info, is_stereo))
    mono = ~is_stereo

    # ---- initial estimates = perturbed truth ---------------------------------
    q_gt = rot_to_quat(Rwc)
    dR = _rodrigues(rng.normal(0, 0.01 / np.sqrt(3), (P, 3)))
    R0 = dR @ Rwc
    t0 = (dR @ twc[..., None])[..., 0] + rng.normal(0, 0.05 / np.sqrt(3), (P, 3))
    q0 = rot_to_quat(R0)
    lm0 = lm_gt + rng.normal(0, 0.10 / np.sqrt(3), (L, 3))
    pose_fixed = np.zeros(P, np.uint8)
    pose_fixed[:fix_first] = 1
    q0[:fix_first] = q_gt[:fix_first]
    t0[:fix_first] = twc[:fix_first]
    lm_fixed = (rng.random(L) < fixed_landmark_frac).astype(np.uint8)
    lm0[lm_fixed == 1] = lm_gt[lm_fixed == 1]

    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    return Graph(
        pose_q=f64(q0), pose_t=f64(t0), pose_fixed=pose_fixed,
        lm_xyz=f64(lm0), lm_fixed=lm_fixed,
        mono_p=i32(pose_of_edge[mono]), mono_l=i32(lm_of_edge[mono]),
        mono_meas=f64(meas[mono, :2]), mono_info=f64(info[mono]),
        stereo_p=i32(pose_of_edge[is_stereo]), stereo_l=i32(lm_of_edge[is_stereo]),
        stereo_meas=f64(meas[is_stereo]), stereo_info=f64(info[is_stereo]),
        camera=f64([fx, fy, cx, cy, bf]), name=name,
        gt_pose_q=f64(q_gt), gt_pose_t=f64(twc), gt_lm_xyz=f64(lm_gt))


def perturb_poses(g: Graph, rot: float, trans: float, seed: int) -> Graph:
    """Left-multiplies every pose estimate by a random rotation (rotation vector ~ N(0, rot^2) rad) and adds
    N(0, trans^2) m to its translation: starts far enough from the optimum that Levenberg-Marquardt REJECTS trials
    (used by the parity cases that pin the reject / lambda * nu path of src/cuda_graph_optimisation.cpp:96-106)."""
    rng = np.random.default_rng(seed)
    rv = rot * rng.standard_normal((g.n_poses, 3))
    th = np.linalg.norm(rv, axis=1, keepdims=True)
    dq = np.concatenate([np.sin(th / 2) * rv / th, np.cos(th / 2)], 1)
    x1, y1, z1, w1 = dq.T
    x2, y2, z2, w2 = g.pose_q.T
    g.pose_q = np.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                         w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], 1)
    g.pose_t = g.pose_t + trans * rng.standard_normal(g.pose_t.shape)
    return g


# ---------------------------------------------------------------------------
# depth edges, pose-only ICP edges, graphs without free landmarks.  These helpers work on PACKED
# problems (cugo_b200.pack.Problem): the reference JSON schema has no records for these edge kinds.
# ---------------------------------------------------------------------------
def stereo_to_depth(prob):
    """Every stereo set (u, v, u_r) becomes a DepthEdgeSet (include/ba_types.h:171-234) with the measurement
    (u, v, 1/z), 1/z = (u - u_r) / bf -- the same observations in the depth camera model."""
    import copy
    out = copy.copy(prob)
    out.sets = []
    for s in prob.sets:
        if s.dim == 3 and getattr(s, "kind", "") == "":
            m = s.meas.copy()
            bf = s.cam[:, 4] if s.cam.shape[0] > 1 else s.cam[0, 4]
            m[:, 2] = (s.meas[:, 0] - s.meas[:, 2]) / bf
            # information of the inverse depth row is shared with the pixel rows (one omega per edge): scale it into pixels
            s = dataclasses.replace(s, meas=np.ascontiguousarray(m), kind="depth")
        out.sets.append(s)
    return out


def fix_all_landmarks(prob):
    """The same problem with every landmark fixed: the reference's doSchur_ == false path (src/block_solver.cpp:80)
    -- motion-only bundle adjustment."""
    import copy
    out = copy.copy(prob)
    out.nL_act = 0
    keep = []
    for s in prob.sets:
        if getattr(s, "kind", "") in ("plane", "line"):
            keep.append(s)
            continue
        act = s.iP < prob.nP_act            # edges between a fixed pose and a (now) fixed landmark are inactive
        idx = np.flatnonzero(act)
        keep.append(dataclasses.replace(
            s, iP=np.ascontiguousarray(s.iP[idx]), iL=np.ascontiguousarray(s.iL[idx]),
            meas=np.ascontiguousarray(s.meas[idx]), src=np.ascontiguousarray(s.src[idx]),
            omega=np.ascontiguousarray(s.omega[idx]) if s.omega.shape[0] > 1 else s.omega,
            cam=np.ascontiguousarray(s.cam[idx]) if s.cam.shape[0] > 1 else s.cam))
    out.sets = keep
    return out


def make_icp_problem(n_poses=6, n_planes=400, n_lines=300, seed=11, noise=0.01, rot=0.05, trans=0.2, n_fixed=0,
                     true_poses=None):
    """Pose-only ICP problem (reference include/icp_types.h): every pose T maps platform points into the world,
    Pw = R(q) p + t.  Scene: random planes (unit normal n, offset d) and line segments (a, b); a match holds the point
    p = T_true^-1 (world point on the plane / line) + noise.  Estimates start from T_true perturbed by rot / trans.
    true_poses ([P,7], optional) replaces the random T_true.
    Returns a packed Problem with a "plane" and a "line" edge set (either may be empty -> left out)."""
    from .pack import Problem, EdgeSetArrays
    rng = np.random.default_rng(seed)
    P = n_poses
    Rt = _rodrigues(0.6 * rng.standard_normal((P, 3)))
    tt = 5.0 * rng.standard_normal((P, 3))
    if true_poses is not None:
        Rt, tt = quat_to_rot(np.asarray(true_poses)[:P, :4]), np.asarray(true_poses)[:P, 4:].copy()
    nact = P - n_fixed
    dR = _rodrigues(rot * rng.standard_normal((P, 3)))
    R0 = dR @ Rt
    t0 = tt + trans * rng.standard_normal((P, 3))
    R0[nact:], t0[nact:] = Rt[nact:], tt[nact:]          # fixed poses sit at the truth
    poses = np.concatenate([rot_to_quat(R0), t0], 1)
    sets = []

    def to_platform(ip, Xw):
        return np.einsum("nji,nj->ni", Rt[ip], Xw - tt[ip]) + noise * rng.standard_normal(Xw.shape)

    if n_planes > 0:
        ip = rng.integers(0, nact, n_planes).astype(np.int32)
        n = rng.standard_normal((n_planes, 3))
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        dist = 3.0 * rng.standard_normal(n_planes)
        X = 4.0 * rng.standard_normal((n_planes, 3))
        X = X - ((X * n).sum(1) - dist)[:, None] * n      # onto the plane
        m = np.concatenate([n, dist[:, None], to_platform(ip, X)], 1)
        sets.append(EdgeSetArrays(dim=1, iP=ip, iL=np.zeros(0, np.int32), meas=np.ascontiguousarray(m),
                                  omega=1.0 + rng.random(n_planes), cam=np.zeros((0, 5)),
                                  src=np.arange(n_planes), kind="plane"))
    if n_lines > 0:
        ip = rng.integers(0, nact, n_lines).astype(np.int32)
        a = 4.0 * rng.standard_normal((n_lines, 3))
        b = a + rng.standard_normal((n_lines, 3)) + 0.5
        X = a + rng.random((n_lines, 1)) * (b - a)
        length = np.linalg.norm(a - b, axis=1)
        # a point exactly on the line has no residual gradient (|c| = 0): keep the matches a little off it
        X = X + 0.05 * rng.standard_normal((n_lines, 3))
        m = np.concatenate([a, b, length[:, None], to_platform(ip, X)], 1)
        sets.append(EdgeSetArrays(dim=1, iP=ip, iL=np.zeros(0, np.int32), meas=np.ascontiguousarray(m),
                                  omega=np.array([2.0]), cam=np.zeros((0, 5)), src=np.arange(n_lines), kind="line"))
    return Problem(poses=np.ascontiguousarray(poses), landmarks=np.zeros((0, 3)), nP_act=nact, nL_act=0,
                   pose_ids=np.arange(P), lm_ids=np.zeros(0, np.int64), sets=sets)
