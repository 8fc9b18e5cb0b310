"""Worker of test_two_gpus_match_one (tests/test_gpu_parity.py): launched with torchrun, one rank per GPU.
Runs the sharded LM optimisation through cugo_b200.dist.DistBlockSolver and, on rank 0, dumps the chi2
trajectory and the gathered estimates."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))


def main():
    import torch
    import torch.distributed as dist
    from cugo_b200 import dist as cdist, pack, synth
    name, niter, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    robust = sys.argv[4] if len(sys.argv) > 4 else "none"
    delta = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
    dist_factor = int(sys.argv[6]) if len(sys.argv) > 6 else 1     # 2: force the subtree-distributed factorisation
    peer_reduce = int(sys.argv[7]) if len(sys.argv) > 7 else 1     # 0: exchange step through ncclAllReduce instead of peer memory
    dev = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl")
    try:
        depth = name.endswith("+depth")          # row f4: the stereo set read as a DepthEdgeSet
        prob = pack.pack(synth.make_graph(name.split("+")[0]))
        if depth:
            prob = synth.stereo_to_depth(prob)
        s = cdist.DistBlockSolver(prob, dev, robust=robust, delta=delta)
        s.s.set_option("dist_factor", dist_factor)
        s.s.set_option("peer_reduce", peer_reduce)
        chi, lam, tr = s.optimize(niter)
        chi2, _, _ = (None, None, None)
        poses, lms = s.estimates()
        # determinism: same inputs, same bits
        s.set_estimates(prob.poses, prob.landmarks)
        chi_b, _, _ = s.optimize(niter)
        if dist.get_rank() == 0:
            np.savez(out, chi2=chi, lam=lam, trials=tr, poses=poses, lms=lms, chi2_again=chi_b,
                     sizes=np.array(list(s.sizes().values())))
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
