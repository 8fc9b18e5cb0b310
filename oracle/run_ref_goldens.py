"""Runs the UNMODIFIED reference (oracle/_ref/ref_driver, built by oracle/Makefile from
/root/reference) on the seeded synthetic graphs ON THE GPU BOX and condenses its dumps into the
golden fixtures the tests pin the oracle and the CUDA engine against (tests/golden/ref_<name>.npz
+ .json).  TEST INFRASTRUCTURE.  Usage (under gpurun):
    python oracle/run_ref_goldens.py gpurun_out/golden tiny small loop kitti07 kitti00
"""
import hashlib
import json
import os
import re
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
from cugo_b200 import graphio, synth  # noqa: E402

DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
INT_ARRAYS = ("hpl_colptr", "hpl_rowind", "hsc_rowptr", "hsc_colind", "csr_rowptr", "csr_colind", "bsr2csr",
              "mulids", "edge2hpl", "mono_edge2pl", "stereo_edge2pl")
F64_ARRAYS = ("Hpp", "Hll", "Hpl", "b", "bsc", "Hsc", "x", "updated_estimates", "final_estimates")
FULL_LIMIT = 4000      # graphs with at most this many edges keep full arrays

# name -> (generator kwargs, robust kernel type (reference enum: 0 none 1 cauchy 2 tukey), delta)
CONFIGS = {
    "tiny": ({}, 0, 1.0), "small": ({}, 0, 1.0), "loop": ({}, 0, 1.0), "kitti07": ({}, 0, 1.0),
    "kitti00": ({}, 0, 1.0),
    "small_cauchy": (dict(outlier_frac=0.05, seed=7002), 1, 2.0),
    "small_tukey": (dict(outlier_frac=0.05, seed=7002), 2, 4.0),
    "small_fixedlm": (dict(fixed_landmark_frac=0.15, seed=21), 0, 1.0),
    # poses start far off (synth.perturb_poses): several outer iterations REJECT trials
    # (src/cuda_graph_optimisation.cpp:96-106: lambda *= nu, nu *= 2, restoreDiagonal, pop)
    "tiny_reject": (dict(seed=11), 0, 1.0),
    # 5 % gross outliers, EdgeSet::setOutlierThreshold(OUTLIER_THR): optimize(5), updateEdges, optimize(5) again
    "small_outlier": (dict(outlier_frac=0.05, seed=7002), 0, 1.0),
}
POSE_NOISE = {"tiny_reject": (0.1, 3.0, 2)}   # synth.perturb_poses(rot [rad], trans [m], seed) before the run
OUTLIER = {"small_outlier": (30.0, 5, 5)}  # threshold on rho(omega |e|^2), niter of the first and second optimize()


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out_dir = sys.argv[1]
    names = sys.argv[2:] or list(CONFIGS)
    os.makedirs(out_dir, exist_ok=True)
    for name in names:
        kw, rk, delta = CONFIGS[name]
        base = name.split("_")[0]
        g = synth.make_graph(base, **kw)
        if name in POSE_NOISE:
            synth.perturb_poses(g, *POSE_NOISE[name])
        bag = os.path.join(out_dir, name + ".bag")
        graphio.write_graph(g, bag)
        prefix = os.path.join(out_dir, "raw_" + name)
        reps = 4
        if name in OUTLIER:
            thr, n1, n2 = OUTLIER[name]
            r = subprocess.run([DRIVER, bag, prefix, str(n1), str(rk), str(delta), "1", repr(thr), str(n2)],
                               capture_output=True, text=True)
            os.remove(bag)
            if r.returncode != 0 or not os.path.exists(prefix + ".json"):
                print(name, "reference driver FAILED rc", r.returncode, r.stderr[-2000:], flush=True)
                continue
            meta = json.load(open(prefix + ".json"))
            arrays = {}
            for a, t in (("mono_edge2pl", np.int32), ("stereo_edge2pl", np.int32), ("mono_outliers_a", np.int32),
                         ("stereo_outliers_a", np.int32), ("mono_outliers_b", np.int32), ("stereo_outliers_b", np.int32),
                         ("final_estimates", np.float64)):
                fn = "%s.%s.bin" % (prefix, a)
                arrays[a] = np.fromfile(fn, t)
                os.remove(fn)
            meta["generator"] = dict(name=base, kwargs=kw)
            json.dump(meta, open(os.path.join(out_dir, "ref_%s.json" % name), "w"), indent=1)
            np.savez_compressed(os.path.join(out_dir, "ref_%s.npz" % name), **arrays)
            print(name, "outliers", meta["outliers_a"], meta["outliers_b"], "chi2_b", meta["chi2_b"][:3], flush=True)
            continue
        r = subprocess.run([DRIVER, bag, prefix, "10", str(rk), str(delta), str(reps)], capture_output=True, text=True)
        open(prefix + ".stdout", "w").write(r.stdout + "\n--- stderr ---\n" + r.stderr)
        os.remove(bag)
        if r.returncode != 0 or not os.path.exists(prefix + ".json"):
            print(name, "reference driver FAILED rc", r.returncode, r.stderr[-2000:], flush=True)
            continue
        meta = json.load(open(prefix + ".json"))
        # rejected trials per outer iteration of the cold run: the reference's verbose line
        # (src/cuda_graph_optimisation.cpp:114-132, "levenberg iterations = q")
        meta["trials"] = [int(v) for v in re.findall(r"levenberg iterations = (\d+)", r.stdout)][:len(meta["chi2"])]
        if name in POSE_NOISE:
            meta["pose_noise"] = list(POSE_NOISE[name])
        arrays, digest = {}, {}
        full = g.n_edges <= FULL_LIMIT
        for a in INT_ARRAYS + F64_ARRAYS:
            fn = "%s.%s.bin" % (prefix, a)
            v = np.fromfile(fn, np.int32 if a in INT_ARRAYS else np.float64)
            os.remove(fn)
            digest[a] = dict(n=int(v.size), sha256=sha(v))
            if full:
                arrays[a] = v
            elif a in F64_ARRAYS:
                arrays[a + "_sub"] = v[::97].copy()       # strided sample of the large value arrays
        meta["arrays"] = digest
        meta["full"] = full
        meta["generator"] = dict(name=base, kwargs=kw)
        json.dump(meta, open(os.path.join(out_dir, "ref_%s.json" % name), "w"), indent=1)
        np.savez_compressed(os.path.join(out_dir, "ref_%s.npz" % name), **arrays)
        print(name, "chi2", meta["chi2"][:3], "...", "opt ms", meta["wall_ms_optimize"], flush=True)


if __name__ == "__main__":
    main()
