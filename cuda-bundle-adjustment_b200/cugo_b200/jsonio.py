"""Reader / writer of the reference's JSON graph schema, without OpenCV.

The reference's sample parses this file with cv::FileStorage
(samples/sample_ba_from_file/main.cpp:78-160); the fields are

  fx fy cx cy bf                                              top-level numbers (one shared camera)
  pose_vertices      [{id, fixed, q: [x,y,z,w], t: [x,y,z]}]
  landmark_vertices  [{id, fixed, Xw: [x,y,z]}]
  monocular_edges    [{vertexP, vertexL, measurement: [u,v],    information}]
  stereo_edges       [{vertexP, vertexL, measurement: [u,v,ur], information}]

Vertex ids are arbitrary integers in the file (ORB-SLAM keyframe / map-point ids).  `read_json` maps them to the
dense 0..n-1 positions of `synth.Graph` in ASCENDING id order -- the order VertexSet::generateEstimateData walks
its std::map (src/optimisable_graph.hpp:84-126) -- and returns the original ids beside the graph so results can be
written back under them.  cv::FileStorage writes an optional "%YAML:1.0"-less plain JSON document, so the standard
library parser reads the reference's own dumps (ba_kitti_*.json, absent from the snapshot) unchanged.
"""
from __future__ import annotations

import json

import numpy as np

from .synth import Graph


def read_json(path: str):
    """-> (Graph, pose_ids [P] int64, landmark_ids [L] int64); ids ascending."""
    with open(path, "r") as f:
        doc = json.load(f)
    cam = np.array([float(doc[k]) for k in ("fx", "fy", "cx", "cy", "bf")], np.float64)

    def vertices(key, fields):
        recs = doc.get(key, [])
        ids = np.array([int(r["id"]) for r in recs], np.int64)
        order = np.argsort(ids, kind="stable")
        if ids.shape[0] and np.any(np.diff(ids[order]) == 0):
            raise ValueError("%s: duplicate vertex id" % key)
        fixed = np.array([int(r["fixed"]) != 0 for r in recs], np.uint8)[order]
        cols = [np.array([[float(v) for v in r[name]][:w] for r in recs], np.float64).reshape(len(recs), w)[order]
                for name, w in fields]
        return ids[order], fixed, cols

    pid, pfix, (q, t) = vertices("pose_vertices", (("q", 4), ("t", 3)))
    lid, lfix, (xw,) = vertices("landmark_vertices", (("Xw", 3),))

    def edges(key, dim):
        recs = doc.get(key, [])
        n = len(recs)
        vp = np.array([int(r["vertexP"]) for r in recs], np.int64)
        vl = np.array([int(r["vertexL"]) for r in recs], np.int64)
        ip, il = np.searchsorted(pid, vp), np.searchsorted(lid, vl)
        if n and (np.any(ip >= pid.shape[0]) or np.any(pid[np.minimum(ip, pid.shape[0] - 1)] != vp) or
                  np.any(il >= lid.shape[0]) or np.any(lid[np.minimum(il, lid.shape[0] - 1)] != vl)):
            raise ValueError("%s: edge refers to an unknown vertex id" % key)
        meas = np.array([[float(v) for v in r["measurement"]][:dim] for r in recs], np.float64).reshape(n, dim)
        info = np.array([float(r["information"]) for r in recs], np.float64)
        return ip.astype(np.int32), il.astype(np.int32), meas, info

    mp, ml, mm, mi = edges("monocular_edges", 2)
    sp, sl, sm, si = edges("stereo_edges", 3)
    g = Graph(pose_q=q, pose_t=t, pose_fixed=pfix, lm_xyz=xw, lm_fixed=lfix, mono_p=mp, mono_l=ml, mono_meas=mm,
              mono_info=mi, stereo_p=sp, stereo_l=sl, stereo_meas=sm, stereo_info=si, camera=cam)
    return g, pid, lid


def write_json(g: Graph, path: str, pose_ids=None, landmark_ids=None) -> None:
    """Writes `g` in the reference's schema; ids default to the dense positions.  Numbers are written with
    17 significant digits so a read-back is bit-exact."""
    pid = np.arange(g.n_poses) if pose_ids is None else np.asarray(pose_ids)
    lid = np.arange(g.n_landmarks) if landmark_ids is None else np.asarray(landmark_ids)
    fl = lambda a: "[" + ", ".join(repr(float(v)) for v in a) + "]"
    with open(path, "w") as f:
        f.write("{\n")
        for k, v in zip(("fx", "fy", "cx", "cy", "bf"), g.camera):
            f.write('"%s": %r,\n' % (k, float(v)))
        f.write('"pose_vertices": [\n')
        f.write(",\n".join('{"id": %d, "fixed": %d, "q": %s, "t": %s}' % (pid[i], int(g.pose_fixed[i]), fl(g.pose_q[i]), fl(g.pose_t[i]))
                           for i in range(g.n_poses)))
        f.write('\n],\n"landmark_vertices": [\n')
        f.write(",\n".join('{"id": %d, "fixed": %d, "Xw": %s}' % (lid[i], int(g.lm_fixed[i]), fl(g.lm_xyz[i]))
                           for i in range(g.n_landmarks)))
        for key, ep, el, meas, info in (("monocular_edges", g.mono_p, g.mono_l, g.mono_meas, g.mono_info),
                                        ("stereo_edges", g.stereo_p, g.stereo_l, g.stereo_meas, g.stereo_info)):
            f.write('\n],\n"%s": [\n' % key)
            f.write(",\n".join('{"vertexP": %d, "vertexL": %d, "measurement": %s, "information": %r}' %
                               (pid[ep[i]], lid[el[i]], fl(meas[i]), float(info[i])) for i in range(ep.shape[0])))
        f.write("\n]\n}\n")
