#!/usr/bin/env python
"""A/B of engine options on one workload:  python scripts/ab_opts.py kitti00 panel_rb=10 panel_rb=5 "panel_rb=5,zinv=0"
Every argument after the workload is one variant (comma-separated name=value pairs, '-' = defaults).  Prints the best
device time of optimize(10) over 5 runs, the final chi2 and the per-kernel event times of one trial summed by kernel."""
import os, sys, collections
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "cuda-bundle-adjustment_b200"))
from cugo_b200 import pack, solver, synth

wl = sys.argv[1]
prob = pack.pack(synth.make_graph(wl))
for variant in sys.argv[2:] or ["-"]:
    s = solver.BlockSolver(0)
    s.initialize(prob)
    if variant != "-":
        for kv in variant.split(","):
            k, v = kv.split("=")
            s.set_option(k, int(v))
    best, chi = 1e9, None
    for rep in range(5):
        s.set_estimates(prob.poses, prob.landmarks)
        st = s.optimize(10)
        best = min(best, s.last_timing()[0])
        chi = st[0][-1] if isinstance(st, tuple) else None
    sums = collections.OrderedDict()
    for name, ms in s.profile_trial():
        sums[name] = sums.get(name, 0.0) + ms
    top = "  ".join("%s %.3f" % (k, v) for k, v in sorted(sums.items(), key=lambda kv: -kv[1])[:6])
    print("%-32s optimize(10) %8.3f ms  chi2 %.10g | %s" % (variant, best, chi if chi is not None else float("nan"), top), flush=True)
    s.close()
