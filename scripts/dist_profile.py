"""Per-kernel event times of one LM trial on every rank of a sharded run (torchrun, one process per GPU).
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/dist_profile.py [workload] [weak|strong]"""
import collections, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
import torch, torch.distributed as dist
from cugo_b200 import pack, synth
from cugo_b200.dist import DistBlockSolver

name = sys.argv[1] if len(sys.argv) > 1 else "kitti00"
mode = sys.argv[2] if len(sys.argv) > 2 else "weak"
dev = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
world, rank = dist.get_world_size(), dist.get_rank()
P0, L0, E0, S0, _, _ = synth.SHAPES[name]
g = synth.make_graph(name, n_landmarks=L0 * world, n_edges=E0 * world, n_stereo=S0 * world) if mode == "weak" else synth.make_graph(name)
prob = pack.pack(g)
s = DistBlockSolver(prob, dev)
s.build_structure()
for rep in range(3):
    s.set_estimates(prob.poses, prob.landmarks)
    prof = s.s.profile_trial()
agg = collections.OrderedDict()
for k, ms in prof:
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += ms
tot = sum(ms for _, ms in prof)
for r in range(world):
    dist.barrier()
    if r == rank:
        print("rank %d: one trial, per kernel (event-timed, serialised): total %.3f ms" % (rank, tot))
        for k, (n, t) in agg.items():
            print("   %-16s x%-3d %8.3f ms" % (k, n, t))
        sys.stdout.flush()
best = 1e9
for rep in range(5):
    s.set_estimates(prob.poses, prob.landmarks)
    s.optimize(10)
    best = min(best, s.last_timing()[0])
if rank == 0:
    print("optimize(10): %.3f ms" % best)
dist.destroy_process_group()
