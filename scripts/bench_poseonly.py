"""Measurement of the SURVEY.md 8(f) row-4 paths (not the headline: bench.py stays on the KITTI-00 shape).

Workloads, 10 LM iterations each, estimates re-set before every run:
  icp_scan     one pose, 1.0 M plane + 0.5 M line matches   (a LiDAR scan against a map: every edge on one pose)
  icp_window   200 poses, 2.0 M plane + 1.0 M line matches  (a window of scans)
  motion_only  the KITTI-00 shape with every landmark fixed  (561 k mono / stereo edges, 1332 poses; the reference's
               doSchur_ == false path)
Per workload: device time of optimize(10) (CUDA events around the call, median of `reps`), the per-kernel times of one
trial (events between the kernels, outside the graph), the oracle's time for the same optimize(10) on one host core
(seconds at these sizes), chi2 agreement, and for the dominant kernel the
algorithmic bytes / time against the measured HBM peak (MEASURED_PEAKS.json, else 6 550 GB/s as the profiling guide says).
Algorithmic bytes of an ICP edge: its match (plane 7, line 10 doubles) + information 8 + set id 1 (+ pose index 4 in
k_icp_errors); motion-only BA: the pose-major record of an edge (uv 16, r 8, omega 8, iL 4, set 1) + its landmark (24).
One JSON line per workload.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6550.0, "fallback (B200_PROFILING.md)"


def run(name, prob, cpu_prob, cpu_scale, edge_bytes, dominant, reps=7, niter=10):
    import oracle
    from cugo_b200 import solver
    solver.BlockSolver(0).close()           # context + library warm
    c0 = time.perf_counter()
    s = solver.BlockSolver(0)
    s.initialize(prob)                      # host copies of the caller's arrays
    c1 = time.perf_counter()
    s.build_structure()                     # H2D (pageable host copies), pose order / slices (ICP), edge orders (BA), trial graph
    c2 = time.perf_counter()
    s.optimize(1)
    ms, chi = [], None
    for _ in range(reps):
        s.set_estimates(prob.poses, prob.landmarks)
        chi, _, tr = s.optimize(niter)
        ms.append(s.last_timing()[0])
    kernels = s.last_timing()[1]
    s.set_estimates(prob.poses, prob.landmarks)
    prof = s.profile_trial()
    agg = {}
    for k, v in prof:
        agg[k] = agg.get(k, 0.0) + v
    o = oracle.Oracle(cpu_prob)
    t0 = time.perf_counter()
    chi_o, _, tr_o = o.optimize(niter)
    cpu_s = (time.perf_counter() - t0) * cpu_scale
    rel = None
    if cpu_prob is prob:     # (a converged run may stop an iteration apart: the last gain ratios are rounding noise)
        k = min(len(chi), len(chi_o))
        rel = float(np.max(np.abs(chi[:k] - chi_o[:k]) / np.abs(chi_o[:k])))
    E = sum(st.n for st in prob.sets)
    med = float(np.median(ms))
    peak, how = hbm_peak()
    dom_ms = agg.get(dominant, 0.0)
    ach = edge_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
    print(json.dumps({
        "workload": name, "edges": E, "poses": int(prob.nP_act), "lm_iterations": len(chi), "lm_iterations_oracle": len(chi_o), "trials": int(len(chi) + tr.sum()),
        "ms_optimize": med, "ms_min": float(np.min(ms)),
        "cold_ms": {"initialize": (c1 - c0) * 1e3, "build_structure": (c2 - c1) * 1e3}, "edges_per_s": E * len(chi) / (med * 1e-3),
        "kernels_launched": int(kernels), "trial_kernels_ms": {k: round(v, 4) for k, v in agg.items()},
        "roofline": {"kernel": dominant, "bound": "hbm", "algorithmic_bytes": edge_bytes, "achieved": ach, "peak": peak,
                     "unit": "GB/s", "frac": (ach / peak) if ach else None, "peak_from": how},
        "cpu_baseline": {"kind": "port", "cores": 1, "s_optimize": cpu_s,
                         "sample": "same problem" if cpu_scale == 1 else "1/%d of the matches, time scaled" % cpu_scale},
        "speedup_vs_cpu_port": cpu_s / (med * 1e-3), "chi2_final": float(chi[-1]), "chi2_rel_vs_oracle": rel}))
    sys.stdout.flush()


def main():
    from cugo_b200 import pack, synth
    only = sys.argv[1:] or ["icp_scan", "icp_window", "motion_only"]
    if "icp_scan" in only:
        p = synth.make_icp_problem(n_poses=1, n_planes=1000000, n_lines=500000, rot=0.05, trans=0.3, seed=1)
        run("icp_scan", p, p, 1, 1000000 * (56 + 9) + 500000 * (80 + 9), "icp_build")
    if "icp_window" in only:
        p = synth.make_icp_problem(n_poses=200, n_planes=2000000, n_lines=1000000, rot=0.05, trans=0.3, seed=2)
        run("icp_window", p, p, 1, 2000000 * (56 + 9) + 1000000 * (80 + 9), "icp_build")
    if "motion_only" in only:
        p = synth.fix_all_landmarks(pack.pack(synth.make_graph("kitti00")))
        E = sum(st.n for st in p.sets)
        run("motion_only_kitti00", p, p, 1, E * (16 + 8 + 8 + 4 + 1 + 24), "build_pose")


if __name__ == "__main__":
    main()
