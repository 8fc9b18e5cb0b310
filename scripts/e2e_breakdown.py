"""Wall-clock breakdown of one end-to-end step through the C ABI (host buffers): which call costs what."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
import torch
from cugo_b200 import pack, solver, synth
name = sys.argv[1] if len(sys.argv) > 1 else "kitti00"
g = synth.make_graph(name); prob = pack.pack(g)
s = solver.BlockSolver(0); s.initialize(prob); s.build_structure(); s.optimize(1)
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hp, hl = pin(prob.poses), pin(prob.landmarks)
hs = [(pin(st.meas), pin(st.omega)) for st in prob.sets]
acc = {}
def t(name, f):
    t0 = time.perf_counter(); r = f(); acc.setdefault(name, []).append((time.perf_counter() - t0) * 1e3); return r
for it in range(8):
    t("set_estimates", lambda: s.set_estimates(hp, hl))
    for i, (m, w) in enumerate(hs):
        t("update_measurements[%d]" % i, lambda: s.update_measurements(i, m, w))
    t("optimize(10)", lambda: s.optimize(10))
    t("estimates", lambda: s.estimates())
    t("last_timing", lambda: s.last_timing())
for k, v in acc.items():
    print("%-26s median %.3f ms  (min %.3f)" % (k, float(np.median(v[2:])), min(v[2:])))
print("device time of optimize: %.3f ms" % s.last_timing()[0])
