"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: landmark sharding (cugo_b200/dist.py) and the
algebra behind the one exchange step -- the ranks' partial reduced camera systems sum to the full one."""
import os
import socket

import numpy as np
import pytest

import dist_worker


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_landmark_ranges_balance():
    from cugo_b200 import dist as cdist
    rng = np.random.default_rng(0)
    w = rng.integers(1, 15, 1000)
    for world in (1, 2, 3, 8):
        b = cdist.landmark_ranges(w, world)
        assert b[0] == 0 and b[-1] == 1000 and np.all(np.diff(b) >= 0)
        loads = [w[b[i]:b[i + 1]].sum() for i in range(world)]
        assert max(loads) - min(loads) <= 2 * w.max()
    assert cdist.landmark_ranges(np.zeros(0, np.int64), 4).tolist() == [0] * 5


@pytest.mark.parametrize("mode", ["strided", "ranges"])
@pytest.mark.parametrize("world", [2, 3])
def test_shards_partition_the_problem(graphs, world, mode):
    from cugo_b200 import dist as cdist
    g, prob = graphs("small", fixed_landmark_frac=0.1, seed=5)
    seen = np.zeros(prob.landmarks.shape[0], int)
    n_edges = 0
    for r in range(world):
        sh = cdist.shard_problem(prob, world, r, mode=mode)
        seen[sh.lm_index] += 1
        n_edges += sh.prob.n_edges
        nL = sh.prob.nL_act
        assert np.all(sh.lm_index[:nL] < prob.nL_act) and np.all(sh.lm_index[nL:] >= prob.nL_act)   # free first, then fixed
        for s in sh.prob.sets:
            assert s.n == 0 or (s.iL.min() >= 0 and s.iL.max() < sh.lm_index.shape[0])
        np.testing.assert_array_equal(sh.prob.poses, prob.poses)                                    # poses replicated
    assert np.all(seen == 1) and n_edges == prob.n_edges


def test_partial_reduced_systems_sum_to_the_full_one(graphs, tmp_path):
    import torch.multiprocessing as mp
    import oracle as orc
    world, name, lam = 2, "small", 50.0
    mp.spawn(dist_worker.run, args=(world, _free_port(), name, lam, str(tmp_path)), nprocs=world, join=True)
    red = np.load(os.path.join(str(tmp_path), "reduced.npz"))
    g, prob = graphs(name)
    P = prob.nP_act
    o = orc.Oracle(prob)
    o.build_structure()
    chi = o.compute_errors()
    o.build_system()
    o.set_lambda(lam)
    assert o.solve() and bool(red["ok"])
    H = dist_worker.dense_from_bsr(o, P)
    n = 36 * P * P
    Hsum = red["sum"][:n].reshape(6 * P, 6 * P) - (world - 1) * lam * np.eye(6 * P)   # every oracle rank added lambda
    scale = np.abs(H).max()
    assert np.abs(Hsum - H).max() < 1e-11 * scale
    bsc = o.values("bsc")
    assert np.abs(red["sum"][n:n + 6 * P] - bsc).max() < 1e-11 * np.abs(bsc).max()
    assert red["sum"][n + 6 * P] == pytest.approx(chi, rel=1e-13)
    assert red["sum"][n + 6 * P + 1] == prob.n_edges and red["sum"][n + 6 * P + 2] == prob.nL_act
    assert np.all(red["cover"] == 1)
    np.testing.assert_array_equal(red["gathered"], prob.landmarks)
    # the reduced system solves for the same pose step as the full oracle
    Hfull = np.triu(Hsum) + np.triu(Hsum, 1).T
    xp = np.linalg.solve(Hfull, red["sum"][n:n + 6 * P])
    assert np.abs(xp - o.values("x")[:6 * P]).max() < 1e-8 * np.abs(xp).max()


def test_shards_keep_the_edge_kind_and_refuse_pose_only_sets(graphs):
    """Row f4 on several GPUs: a depth set stays a depth set on every shard (and the shards' chi2 add up to the whole graph's
    in the oracle); pose-only ICP sets have no landmarks to shard by and are refused."""
    import oracle
    from cugo_b200 import dist as cdist, synth
    prob = synth.stereo_to_depth(graphs("small")[1])
    total = oracle.Oracle(prob)
    total.build_structure()
    parts = 0.0
    for r in range(2):
        sh = cdist.shard_problem(prob, 2, r)
        assert [getattr(s, "kind", "") for s in sh.prob.sets] == [getattr(s, "kind", "") for s in prob.sets]
        o = oracle.Oracle(sh.prob)
        o.build_structure()
        parts += o.compute_errors()
    assert parts == pytest.approx(total.compute_errors(), rel=1e-12)
    with pytest.raises(ValueError):
        cdist.shard_problem(synth.make_icp_problem(), 2, 0)
