"""Worker of tests/test_dist_cpu.py: one gloo rank.  TEST CODE: uses the oracle as the per-rank checker of the
sharding algebra (the product's per-rank work is the CUDA engine; its collective is NCCL)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "cuda-bundle-adjustment_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def dense_from_bsr(o, P):
    """Dense upper-triangular 6P x 6P matrix from the oracle's Hsc BSR (blocks column-major)."""
    rp, ci = o.structure("hsc_rowptr"), o.structure("hsc_colind")
    vals = o.values("Hsc").reshape(-1, 6, 6)
    H = np.zeros((6 * P, 6 * P))
    for r in range(P):
        for k in range(rp[r], rp[r + 1]):
            c = ci[k]
            H[6 * r:6 * r + 6, 6 * c:6 * c + 6] = vals[k].T
    return H


def run(rank, world, port, name, lam, out_dir):
    import torch
    import torch.distributed as dist
    import oracle as orc
    from cugo_b200 import dist as cdist, pack, synth
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        prob = pack.pack(synth.make_graph(name))
        sh = cdist.shard_problem(prob, world, rank)
        P = prob.nP_act
        o = orc.Oracle(sh.prob)
        o.build_structure()
        chi = o.compute_errors()
        o.build_system()
        maxd = o.max_diagonal()
        o.set_lambda(lam)
        ok = o.solve()                       # forms the rank's partial Hsc / bsc (and factorises it: PD thanks to lambda)
        H = dense_from_bsr(o, P)
        bsc = o.values("bsc")
        t = torch.from_numpy(np.concatenate([H.reshape(-1), bsc, [chi, float(sh.prob.n_edges), float(sh.prob.nL_act)]]))
        dist.all_reduce(t)                    # the exchange step (NCCL in the product, gloo here)
        tm = torch.tensor([maxd], dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        # landmark estimates gathered back through the same index maps the product uses
        parts = [None] * world
        dist.all_gather_object(parts, (sh.lm_index, sh.prob.landmarks))
        if rank == 0:
            full = np.zeros_like(prob.landmarks)
            cover = np.zeros(prob.landmarks.shape[0], np.int64)
            for idx, val in parts:
                full[idx] = val
                cover[idx] += 1
            np.savez(os.path.join(out_dir, "reduced.npz"), sum=t.numpy(), ok=ok, gathered=full, cover=cover,
                     local_maxdiag=tm.numpy())
    finally:
        dist.destroy_process_group()
