"""Multi-GPU host plumbing: one process per GPU, `torch.distributed` for rendezvous only.

The path shards by LANDMARK: rank r owns a contiguous range of the landmark index
space (balanced by edge count) together with every edge of those landmarks; poses are replicated.  All
per-edge / per-landmark stages are then local.  The exchange step of an LM trial -- the sum of the ranks'
partial reduced camera systems [Hsc | bsc] -- is an ncclAllReduce issued by the engine itself on its
stream (csrc/engine.cu), inside the captured trial graph; this module only
  * cuts the packed problem into shards (`shard_problem`),
  * hands every rank's engine the NCCL unique id (`DistBlockSolver`, broadcast over the process group),
  * gathers the landmark estimates back (`DistBlockSolver.estimates`).
The reference is single-GPU; parity is checked against the 1-GPU result (tests/test_dist_cpu.py checks the
sharding algebra on CPU with gloo, tests/test_gpu_parity.py::test_two_gpus_match_one on a 2-GPU box).
"""
from __future__ import annotations

import dataclasses

import numpy as np

from .pack import EdgeSetArrays, Problem


@dataclasses.dataclass
class Shard:
    prob: Problem              # the rank's problem: all poses, local landmarks (free first, then fixed)
    lm_index: np.ndarray       # [L_local] global landmark index of every local landmark
    rank: int
    world: int


def landmark_ranges(weights: np.ndarray, world: int) -> np.ndarray:
    """Cut 0..n into `world` contiguous ranges of (almost) equal total weight: bounds[world + 1]."""
    n = weights.shape[0]
    if n == 0:
        return np.zeros(world + 1, np.int64)
    csum = np.cumsum(weights, dtype=np.int64)
    total = int(csum[-1])
    targets = (np.arange(1, world) * total) // world
    cuts = np.searchsorted(csum, targets, side="left") + 1
    cuts = np.minimum(cuts, n)
    return np.concatenate([[0], np.maximum.accumulate(cuts), [n]]).astype(np.int64)


def shard_problem(prob: Problem, world: int, rank: int, mode: str = "strided") -> Shard:
    """Rank `rank`'s shard of a packed problem (index order: free vertices first, then fixed ones).

    mode "strided" (default): landmark l belongs to rank l % world.  Neighbouring landmarks are seen by the same
    poses, so every shard is a thinned copy of the whole graph: the same poses, pose degrees and Hsc block pattern
    as one GPU would have at 1/world of the landmarks -- the per-pose work units of the build and Schur kernels
    stay balanced.  mode "ranges": contiguous landmark ranges balanced by edge count; a range covers a stretch of
    the trajectory, so a shard concentrates world x the edges on 1/world of the poses (measured on 8 GPUs, KITTI-00
    shape: 18.5 ms per optimize(10) against 16.6 ms after slicing the pose-major pass)."""
    L_all, nL = prob.landmarks.shape[0], prob.nL_act
    if world < 1 or not 0 <= rank < world:
        raise ValueError("shard_problem: bad rank %d / world %d" % (rank, world))
    if mode == "strided" and world > 1:
        # Every rank checks EVERY rank's shard, so all ranks fail together here instead of one rank throwing in
        # build_structure (no free landmarks / no edges) while its peers block in the first collective.
        owner_edges = np.zeros(world, np.int64)
        for s in prob.sets:
            owner = np.where(s.iL < nL, s.iL % world, (s.iL - nL) % world)
            owner_edges += np.bincount(owner, minlength=world)
        if nL < world or np.any(owner_edges == 0):
            raise ValueError("shard_problem: %d free landmarks / per-rank edge counts %s cannot feed %d ranks "
                             "(every rank needs a free landmark and an edge)" % (nL, owner_edges.tolist(), world))
    if mode == "strided":
        own_free = np.arange(rank, nL, world)
        own_fixed = nL + np.arange(rank, L_all - nL, world)
    else:
        deg = np.zeros(L_all, np.int64)
        for s in prob.sets:
            np.add.at(deg, s.iL, 1)
        deg = np.maximum(deg, 1)
        fb = landmark_ranges(deg[:nL], world)                  # free landmarks
        xb = landmark_ranges(deg[nL:], world) + nL             # fixed landmarks
        own_free = np.arange(fb[rank], fb[rank + 1])
        own_fixed = np.arange(xb[rank], xb[rank + 1])
    lm_index = np.concatenate([own_free, own_fixed]).astype(np.int64)
    local_of = np.full(L_all, -1, np.int64)
    local_of[lm_index] = np.arange(lm_index.shape[0])
    sets = []
    for s in prob.sets:
        kind = getattr(s, "kind", "")
        if kind in ("plane", "line"):
            raise ValueError("shard_problem: pose-only ICP edge sets have no landmarks to shard by; they run on one GPU")
        keep = np.flatnonzero(local_of[s.iL] >= 0)
        per_edge_w = s.omega.shape[0] == s.n and s.n > 1
        per_edge_c = s.cam.shape[0] == s.n and s.n > 1
        sets.append(EdgeSetArrays(
            dim=s.dim,
            iP=np.ascontiguousarray(s.iP[keep], np.int32),
            iL=np.ascontiguousarray(local_of[s.iL[keep]], np.int32),
            meas=np.ascontiguousarray(s.meas[keep]),
            omega=np.ascontiguousarray(s.omega[keep] if per_edge_w else s.omega[:1]),
            cam=np.ascontiguousarray(s.cam[keep] if per_edge_c else s.cam[:1]),
            src=s.src[keep], kind=kind))
    local = Problem(poses=prob.poses, landmarks=np.ascontiguousarray(prob.landmarks[lm_index]),
                    nP_act=prob.nP_act, nL_act=int(own_free.shape[0]), pose_ids=prob.pose_ids,
                    lm_ids=prob.lm_ids[lm_index], sets=sets)
    return Shard(prob=local, lm_index=lm_index, rank=rank, world=world)


class DistBlockSolver:
    """BlockSolver over `world` GPUs of one node.  Every rank constructs it with the SAME packed problem
    (or with its own shard through `from_shard`); `optimize` runs the device-resident LM loop on all ranks."""

    def __init__(self, prob: Problem, device: int, group=None, robust="none", delta=1.0, shard: Shard | None = None,
                 share_comm=None):
        # share_comm is accepted for call-site clarity only: every solver owns its own NCCL communicator (one per
        # engine stream), created below from a fresh unique id
        import torch
        import torch.distributed as dist
        from . import solver
        self.dist, self.torch, self.group = dist, torch, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.shard = shard if shard is not None else shard_problem(prob, self.world, self.rank)
        self.full_L = prob.landmarks.shape[0] if prob is not None else None
        self.s = solver.BlockSolver(device)
        if self.world > 1:
            uid = [solver.BlockSolver.comm_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(uid, src=0, group=group)
            self.s.comm_init(self.world, self.rank, uid[0])
        self.s.initialize(self.shard.prob, robust=robust, delta=delta)

    def build_structure(self):
        self.s.build_structure()

    def optimize(self, niter):
        return self.s.optimize(niter)

    def set_estimates(self, poses, landmarks_full):
        self.s.set_estimates(poses, np.ascontiguousarray(landmarks_full[self.shard.lm_index]))

    def local_estimates(self):
        return self.s.estimates()

    def estimates(self):
        """(poses, landmarks of the FULL problem in index order), gathered over the ranks."""
        poses, lms = self.s.estimates()
        if self.world == 1:
            return poses, lms
        parts = [None] * self.world
        self.dist.all_gather_object(parts, (self.shard.lm_index, lms), group=self.group)
        full = np.zeros((sum(p[0].shape[0] for p in parts), 3))
        for idx, val in parts:
            full[idx] = val
        return poses, full

    def sizes(self):
        return self.s.sizes()

    def last_timing(self):
        return self.s.last_timing()
