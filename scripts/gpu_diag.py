"""Stage-by-stage diagnostics of the CUDA engine against the oracle (run on the GPU box).
Prints one line per stage and never stops at the first mismatch."""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "cuda-bundle-adjustment_b200"), os.path.join(ROOT, "oracle")]
import oracle as orc  # noqa: E402
from cugo_b200 import pack, solver, synth  # noqa: E402


def rel(a, b):
    if a.shape != b.shape:
        return "SHAPE %s vs %s" % (a.shape, b.shape)
    if a.size == 0:
        return "empty"
    return "%.3e" % (np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def run(name, niter=10, **kw):
    print("=== %s %s" % (name, kw), flush=True)
    g = synth.make_graph(name, **kw)
    prob = pack.pack(g)
    s = solver.BlockSolver(0)
    s.initialize(prob)
    o = orc.Oracle(prob, threads=8)
    t = time.time(); s.build_structure(); print("  gpu build_structure %.1f ms" % ((time.time() - t) * 1e3), s.sizes(), flush=True)
    o.build_structure()
    for k in ("hpl_colptr", "hpl_rowind", "edge2hpl", "hsc_rowptr", "hsc_colind", "csr_rowptr", "csr_colind", "bsr2csr", "mulids"):
        a, b = s.structure(k), o.structure(k)
        print("  struct %-11s %s" % (k, "OK" if a.shape == b.shape and np.array_equal(a, b) else "MISMATCH %s %s" % (a.shape, b.shape)), flush=True)
    print("  chi2  gpu %.12e  oracle %.12e" % (s.compute_errors(), o.compute_errors()), flush=True)
    s.build_system(); o.build_system()
    for k in ("Hpp", "Hll", "Hpl", "b"):
        print("  %-7s relerr %s" % (k, rel(s.values(k), o.values(k))), flush=True)
    md = o.max_diagonal(); print("  maxdiag gpu %.12e oracle %.12e" % (s.max_diagonal(), md))
    lam = 1e-5 * md
    s.set_lambda(lam); o.set_lambda(lam)
    print("  solve ok gpu %s oracle %s" % (s.solve(), o.solve()), flush=True)
    for k in ("invHll", "bsc", "Hsc", "x"):
        print("  %-7s relerr %s" % (k, rel(s.values(k), o.values(k))), flush=True)
    xg, xo = s.values("x"), o.values("x")
    nP6 = 6 * prob.nP_act
    print("  xp relerr %s  xl relerr %s" % (rel(xg[:nP6], xo[:nP6]), rel(xg[nP6:], xo[nP6:])))
    print("  scale gpu %.12e oracle %.12e" % (s.compute_scale(lam), o.compute_scale(lam)))
    s.update(); o.update()
    pg, lg = s.estimates(); po, lo = o.estimates()
    print("  update pose maxabs %.3e lm maxabs %.3e" % (np.abs(pg - po).max(), np.abs(lg - lo).max()))
    print("  chi2 after gpu %.12e oracle %.12e" % (s.compute_errors(), o.compute_errors()), flush=True)
    s.set_estimates(prob.poses, prob.landmarks)
    o = orc.Oracle(prob, threads=8)
    cg, lg_, tg = s.optimize(niter)
    ms, nk = s.last_timing()
    co, lo_, to = o.optimize(niter)
    print("  optimize: %.3f ms device, %d kernels, trials gpu %s oracle %s" % (ms, nk, tg.tolist(), to.tolist()))
    n = min(len(cg), len(co))
    print("  chi2 traj relerr %s" % (np.abs(cg[:n] - co[:n]) / co[:n]).tolist(), flush=True)
    for rep in range(3):
        s.set_estimates(prob.poses, prob.landmarks)
        s.optimize(niter)
        print("  optimize rep %d: %.3f ms" % (rep, s.last_timing()[0]), flush=True)


if __name__ == "__main__":
    for nm in sys.argv[1:] or ["tiny", "small", "loop", "kitti07", "kitti00"]:
        try:
            run(nm)
        except Exception:
            traceback.print_exc()
            print("  FAILED %s" % nm, flush=True)
