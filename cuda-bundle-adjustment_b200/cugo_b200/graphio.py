"""Binary graph file ("BAG1") shared by the Python harness and the C++ programs
(the reference driver used by the parity tests, the facade sample).  It carries exactly the fields of the
reference's JSON schema (samples/sample_ba_from_file/main.cpp:88-160) as raw
little-endian arrays:

  int32  magic 'BAG1', P, L, Em, Es, 0
  f64    camera[5]                      fx fy cx cy bf
  f64    pose[P][7]                     qx qy qz qw tx ty tz   (by vertex id)
  f64    landmark[L][3]
  f64    mono_meas[Em][2], mono_info[Em], stereo_meas[Es][3], stereo_info[Es]
  int32  pose_fixed[P], lm_fixed[L], mono_p[Em], mono_l[Em], stereo_p[Es], stereo_l[Es]
"""
from __future__ import annotations

import numpy as np

from .synth import Graph

MAGIC = 0x31474142


def write_graph(g: Graph, path: str) -> None:
    with open(path, "wb") as f:
        np.array([MAGIC, g.n_poses, g.n_landmarks, g.mono_p.shape[0], g.stereo_p.shape[0], 0],
                 np.int32).tofile(f)
        for a in (g.camera, np.concatenate([g.pose_q, g.pose_t], 1), g.lm_xyz, g.mono_meas,
                  g.mono_info, g.stereo_meas, g.stereo_info):
            np.ascontiguousarray(a, np.float64).tofile(f)
        for a in (g.pose_fixed, g.lm_fixed, g.mono_p, g.mono_l, g.stereo_p, g.stereo_l):
            np.ascontiguousarray(a, np.int32).tofile(f)


def read_graph(path: str) -> Graph:
    with open(path, "rb") as f:
        hdr = np.fromfile(f, np.int32, 6)
        if hdr[0] != MAGIC:
            raise ValueError("not a BAG1 file: %s" % path)
        P, L, Em, Es = (int(v) for v in hdr[1:5])
        rd = lambda n, t=np.float64: np.fromfile(f, t, n)
        cam = rd(5)
        pose = rd(P * 7).reshape(P, 7)
        lm = rd(L * 3).reshape(L, 3)
        mm, mi = rd(Em * 2).reshape(Em, 2), rd(Em)
        sm, si = rd(Es * 3).reshape(Es, 3), rd(Es)
        pf, lf = rd(P, np.int32), rd(L, np.int32)
        mp, ml, sp, sl = rd(Em, np.int32), rd(Em, np.int32), rd(Es, np.int32), rd(Es, np.int32)
    return Graph(pose_q=pose[:, :4].copy(), pose_t=pose[:, 4:].copy(), pose_fixed=pf.astype(np.uint8),
                 lm_xyz=lm, lm_fixed=lf.astype(np.uint8), mono_p=mp, mono_l=ml, mono_meas=mm,
                 mono_info=mi, stereo_p=sp, stereo_l=sl, stereo_meas=sm, stereo_info=si, camera=cam)


ICP_MAGIC = 0x31504349


def write_icp(prob, path: str) -> None:
    """ICP1 file of a packed pose-only problem (synth.make_icp_problem; pose ids = indices), read by
    samples/sample_icp.cpp and the reference driver of the row-f4 goldens:

      int32  magic 'ICP1', P, nPlane, nLine
      f64    pose[P][7]; int32 fixed[P]
      int32  plane_pose[nPlane]; f64 plane_match[nPlane][7] (normal, originDistance, pointP); f64 plane_info[nPlane]
      int32  line_pose[nLine];   f64 line_match[nLine][10] (a, b, length, pointP);            f64 line_info[nLine]
    """
    plane = [s for s in prob.sets if getattr(s, "kind", "") == "plane"]
    line = [s for s in prob.sets if getattr(s, "kind", "") == "line"]
    with open(path, "wb") as f:
        P = prob.poses.shape[0]
        np.array([ICP_MAGIC, P, plane[0].n if plane else 0, line[0].n if line else 0], np.int32).tofile(f)
        np.ascontiguousarray(prob.poses, np.float64).tofile(f)
        (np.arange(P) >= prob.nP_act).astype(np.int32).tofile(f)
        for sets in (plane, line):
            if not sets:
                continue
            s = sets[0]
            om = s.omega if s.omega.shape[0] == s.n else np.full(s.n, s.omega[0])
            np.ascontiguousarray(s.iP, np.int32).tofile(f)
            np.ascontiguousarray(s.meas, np.float64).tofile(f)
            np.ascontiguousarray(om, np.float64).tofile(f)
