"""Where the cold call's host side goes: solver creation, set_vertices, add_edge_set per set, build_structure (wall clock)."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
from cugo_b200 import pack, solver, synth  # noqa: E402

prob = pack.pack(synth.make_graph(sys.argv[1] if len(sys.argv) > 1 else "kitti00"))
w = solver.BlockSolver(0); w.initialize(prob); w.optimize(2)          # context, library, clocks warm
for rep in range(3):
    t = [time.perf_counter()]
    s = solver.BlockSolver(0); t.append(time.perf_counter())
    lib, h = s.lib, s.h
    poses, lms = np.ascontiguousarray(prob.poses), np.ascontiguousarray(prob.landmarks)
    p = lambda a, ty: a.ctypes.data_as(C.POINTER(ty))
    lib.cugo_b200_set_vertices(h, poses.shape[0], prob.nP_act, p(poses, C.c_double), lms.shape[0], prob.nL_act, p(lms, C.c_double))
    t.append(time.perf_counter())
    for st in prob.sets:
        sid = C.c_int()
        lib.cugo_b200_add_edge_set(h, st.dim, st.n, p(st.iP, C.c_int), p(st.iL, C.c_int), p(st.meas, C.c_double), p(st.omega, C.c_double),
                                   int(st.omega.shape[0]), p(st.cam, C.c_double), int(st.cam.shape[0]), 0, C.c_double(1.0), C.byref(sid))
        t.append(time.perf_counter())
    s.prob = prob
    s.build_structure(); t.append(time.perf_counter())
    s.optimize(10); t.append(time.perf_counter())
    names = ["create", "set_vertices"] + ["add_edge_set[%d] n=%d" % (i, st.n) for i, st in enumerate(prob.sets)] + ["build_structure", "optimize(10)"]
    print("rep", rep, "  ".join("%s %.2f ms" % (n, (b - a) * 1e3) for n, a, b in zip(names, t[:-1], t[1:])))
    s.close()
