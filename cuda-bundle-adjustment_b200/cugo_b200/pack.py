"""Host-side packing of an id-keyed graph into the index-space arrays the C ABI
takes.  Mirrors the reference's host packing:

* VertexSet::generateEstimateData (src/optimisable_graph.hpp:84-126): iterate
  vertices in ascending id; non-fixed vertices get indices 0..active-1, fixed
  vertices the trailing indices; estimates stored in that order.
* EdgeSet::init (src/optimisable_graph.hpp:474-572): an edge is kept ("active")
  unless both of its vertices are fixed; per edge (iP, iL), measurement,
  information (omega) and camera are packed.
"""
from __future__ import annotations

import dataclasses
import numpy as np

from .synth import Graph


@dataclasses.dataclass
class EdgeSetArrays:
    dim: int                 # 2 mono (EdgeSE3ProjectXYZ), 3 stereo (EdgeStereoSE3ProjectXYZ)
    iP: np.ndarray           # [n] int32 pose index
    iL: np.ndarray           # [n] int32 landmark index
    meas: np.ndarray         # [n,dim] float64
    omega: np.ndarray        # [n] or [1] float64
    cam: np.ndarray          # [n,5] or [1,5] float64
    src: np.ndarray          # [n] position of the edge in the graph's edge list
    # "" = by dim; "depth" = DepthEdgeSet (dim 3: u, v, 1/z); "plane" / "line" = the pose-only ICP sets (dim 1; iL and cam
    # empty; meas [n,7] normal, originDistance, pointP / [n,10] a, b, length, pointP -- src/measurements.h:13-101)
    kind: str = ""

    @property
    def n(self):
        return int(self.iP.shape[0])


@dataclasses.dataclass
class Problem:
    poses: np.ndarray        # [P_all,7] qx qy qz qw tx ty tz, index order
    landmarks: np.ndarray    # [L_all,3]
    nP_act: int
    nL_act: int
    pose_ids: np.ndarray     # [P_all] id of the vertex at each index
    lm_ids: np.ndarray       # [L_all]
    sets: list               # [EdgeSetArrays]; mono first, then stereo (sample main.cpp:163-164)

    @property
    def n_edges(self):
        return sum(s.n for s in self.sets)


def _index_vertices(fixed: np.ndarray):
    ids = np.arange(fixed.shape[0])
    order = np.concatenate([ids[fixed == 0], ids[fixed != 0]])   # ids ascending inside each part
    index_of_id = np.empty_like(order)
    index_of_id[order] = np.arange(order.shape[0])
    return order, index_of_id, int((fixed == 0).sum())


def pack(g: Graph, per_edge_information=True, per_edge_camera=False) -> Problem:
    p_order, p_index, nP = _index_vertices(g.pose_fixed)
    l_order, l_index, nL = _index_vertices(g.lm_fixed)
    poses = np.ascontiguousarray(np.concatenate([g.pose_q, g.pose_t], axis=1)[p_order])
    lms = np.ascontiguousarray(g.lm_xyz[l_order])
    sets = []
    for dim, ep, el, meas, info in ((2, g.mono_p, g.mono_l, g.mono_meas, g.mono_info),
                                    (3, g.stereo_p, g.stereo_l, g.stereo_meas, g.stereo_info)):
        if ep.shape[0] == 0:
            continue
        active = ~((g.pose_fixed[ep] != 0) & (g.lm_fixed[el] != 0))
        src = np.flatnonzero(active)
        n = src.shape[0]
        if n == 0:
            continue
        omega = info[src] if per_edge_information else info[src][:1]
        cam = np.tile(g.camera, (n, 1)) if per_edge_camera else g.camera.reshape(1, 5)
        sets.append(EdgeSetArrays(
            dim=dim,
            iP=np.ascontiguousarray(p_index[ep[src]], dtype=np.int32),
            iL=np.ascontiguousarray(l_index[el[src]], dtype=np.int32),
            meas=np.ascontiguousarray(meas[src], dtype=np.float64),
            omega=np.ascontiguousarray(omega, dtype=np.float64),
            cam=np.ascontiguousarray(cam, dtype=np.float64),
            src=src))
    return Problem(poses=poses, landmarks=lms, nP_act=nP, nL_act=nL,
                   pose_ids=p_order, lm_ids=l_order, sets=sets)


def unpack_estimates(prob: Problem, poses: np.ndarray, landmarks: np.ndarray):
    """Index-order estimates -> id-order (q[P,4], t[P,3], xyz[L,3]); the
    reference's VertexSet::finalise (src/optimisable_graph.hpp:137-154)."""
    P = np.empty_like(poses)
    P[prob.pose_ids] = poses
    Lm = np.empty_like(landmarks)
    Lm[prob.lm_ids] = landmarks
    return P[:, :4], P[:, 4:], Lm
