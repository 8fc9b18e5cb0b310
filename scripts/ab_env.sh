#!/bin/bash
# A/B of the scheduling variants (cugo_b200_set_option): optimize(10) on kitti00 with each option off, then all off.
python - <<'PY'
import os, sys
sys.path.insert(0, os.path.join(os.getcwd(), "cuda-bundle-adjustment_b200"))
from cugo_b200 import pack, solver, synth
prob = pack.pack(synth.make_graph("kitti00"))
opts = ["pdl", "fuse_build", "panel_dc", "zinv", "bwd_split"]
for off in [[]] + [[o] for o in opts] + [opts]:
    s = solver.BlockSolver(0); s.initialize(prob)
    for o in off: s.set_option(o, 0)
    best = 1e9
    for rep in range(5):
        s.set_estimates(prob.poses, prob.landmarks); s.optimize(10); best = min(best, s.last_timing()[0])
    print("off=%-40s optimize(10) %.3f ms" % (",".join(off) or "-", best)); s.close()
PY
