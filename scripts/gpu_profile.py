"""Per-kernel device times of one LM trial (CUDA events, outside the graph) + optimize() timing."""
import collections
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "cuda-bundle-adjustment_b200")]
from cugo_b200 import pack, solver, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "kitti00"
prob = pack.pack(synth.make_graph(name))
s = solver.BlockSolver(0)
s.initialize(prob)
for kv in filter(None, os.environ.get("CUGO_OPTS", "").split(",")):      # e.g. CUGO_OPTS=panel_rb=5,zinv=0
    s.set_option(kv.split("=")[0], int(kv.split("=")[1]))
s.build_structure()
print(s.sizes())
for rep in range(3):
    s.set_estimates(prob.poses, prob.landmarks)
    prof = s.profile_trial()
agg = collections.OrderedDict()
for k, ms in prof:
    a = agg.setdefault(k, [0, 0.0, 0.0])
    a[0] += 1; a[1] += ms; a[2] = max(a[2], ms)
tot = sum(ms for _, ms in prof)
print("one trial, per kernel (event-timed, launches serialised): total %.3f ms" % tot)
for k, (n, t, mx) in agg.items():
    print("  %-14s x%-3d %8.3f ms  (max %.3f)  %5.1f %%" % (k, n, t, mx, 100 * t / tot))
print("chol per level:")
for nm in ("chol_diag", "chol_trsm", "chol_panel", "chol_update", "chol_bwd"):
    v = np.round([ms for k, ms in prof if k == nm], 3).tolist()
    if v:
        print("  %-12s" % nm, v)
DIAG = False
if os.environ.get("CUGO_B200_DEBUG"):
    import ctypes as C
    a = (C.c_longlong * 8192)()
    DIAG = s.lib.cugo_b200_debug_counters(a, 8192) == 0      # -5: not a diagnostic build (make DIAG=1)
    if not DIAG:
        print("(phase clocks need a diagnostic build of the library: make -C cuda-bundle-adjustment_b200 clean all DIAG=1)")
if DIAG:
    nun = min(2048, int(np.count_nonzero(np.array(a[:2048], dtype=np.int64))))
    ph1 = np.array(a[2048:2048 + nun], dtype=np.int64)
    rnd = np.array(a[4096:4096 + nun], dtype=np.int64); blk = np.array(a[6144:6144 + nun], dtype=np.int64)
    print("schur warp 0 of every unit: rounds mean %.1f max %d; blocks (1000 per shared big block) mean %.1f max %d" % (rnd.mean(), rnd.max(), blk.mean(), blk.max()))
    print("schur units: %d; phase 1 (T stage) cycles per unit  mean %.0f  median %.0f  max %.0f" % (nun, ph1.mean(), np.median(ph1), ph1.max()))
    cyc = np.array(a[:nun], dtype=np.int64)
    print("schur rows: cycles per row  mean %.0f  median %.0f  p99 %.0f  max %.0f  (sum/148/2 = %.0f)" % (
        cyc.mean(), np.median(cyc), np.percentile(cyc, 99), cyc.max(), cyc.sum() / 296))
    print("  top rows", np.argsort(cyc)[-8:].tolist(), np.sort(cyc)[-8:].tolist())
    dc = s.debug_counters()
    print("panel phases of CTA 0 per level (cycles): zero, hsc scatter, child pull, factor, store")
    for lev in range(min(s.sizes()["nlevels"], 64)):
        print("  lev %2d panel" % lev, np.diff(dc[lev, :6]).tolist(), " update (stage, mma, pull+store)",
              np.diff(dc[lev, 8:12]).tolist(), " bwd (load, gemv, trsv)", np.diff(dc[lev, 16:20]).tolist(),
              " factor loop (warp 0 diag update + factor, wait for workers, row solve, whole loop)", dc[lev, 12:16].tolist(), " worker (rank6, look-ahead)", dc[lev, 20:22].tolist())
    trAll = s.debug_counters(176).reshape(-1)
    tr8 = trAll[:128].reshape(16, 8)
    print("traced CTA: R = %d rows, nt = %d block columns; slab warps done at" % (trAll[126], trAll[127]), trAll[128:132].tolist(),
          " row warps done at", trAll[136:144].tolist())
    print("data-flow panel loop, one level, CTA 0, per block row r, cycles since the loop start: chain (pubW = r, preReady[r] seen, pubL = r, "
          "D_r factored) | the row warp's preReady[r] = 1")
    for jb in range(1, 16):
        print("   r %2d" % jb, tr8[jb, :4].tolist(), tr8[jb, 4])
for rep in range(5):
    s.set_estimates(prob.poses, prob.landmarks)
    chi, lam, tr = s.optimize(10)
    ms, nk = s.last_timing()
    print("optimize(10): %.3f ms device, %d kernels, %.3f ms/iter, %.2f M edges/s" % (ms, nk, ms / len(chi), prob.n_edges * len(chi) / ms / 1e3))

if DIAG:
    # timeline of the last trial of the last optimize() (graph, programmatic dependent launch): CTA 0 of every level
    import ctypes as C
    a = (C.c_longlong * 512)()
    s.lib.cugo_b200_debug_counters(a, 512)
    tm = np.array(a[:], dtype=np.int64).reshape(64, 8)
    nlev = s.sizes()["nlevels"]
    t0 = tm[0, 0]
    print("timeline of the last trial, CTA 0 of every level (us since the first panel CTA passed its wait; n/a: that stamp was not "
          "written in this trial, e.g. CTA 0 of a shared backward solve was not the last arriver): "
          "panel [start, end]  update [start, end]  bwd [start, end]")
    def us(a, b):
        dt = (a - b) / 1e3
        return "%.1f" % dt if a > 0 and b > 0 and 0 <= dt < 1e5 else "n/a"
    for lev in range(min(nlev, 64)):
        print("  lev %2d  [%s]  bwd: gather %s  product L21^T x %s  solve+store %s us" % (
            lev, ", ".join(us(tm[lev, k], t0) for k in range(6)),
            us(tm[lev, 7], tm[lev, 4]), us(tm[lev, 6], tm[lev, 7]), us(tm[lev, 5], tm[lev, 6])))
