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
