"""Cold structure build (initialize + build_structure) wall time, per workload."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
import torch
from cugo_b200 import pack, solver, synth
for name in sys.argv[1:] or ["kitti00"]:
    g = synth.make_graph(name); prob = pack.pack(g)
    ts = []
    for rep in range(3):
        s = solver.BlockSolver(0)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        s.initialize(prob); t1 = time.perf_counter()
        s.build_structure(); torch.cuda.synchronize(); t2 = time.perf_counter()
        c, _, _ = s.optimize(10); t3 = time.perf_counter()
        ts.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
        s.close()
    print(name, "initialize / build_structure / optimize(10) ms:", [tuple(round(x, 2) for x in t) for t in ts], "chi2", c[-1])
