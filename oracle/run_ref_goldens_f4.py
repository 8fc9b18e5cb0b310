"""Row f4 goldens: runs the UNMODIFIED reference (oracle/_ref/ref_driver_f4, built by oracle/Makefile from /root/reference)
ON THE GPU BOX on a depth graph and a pose-only ICP problem and condenses its dumps into tests/golden/ref_f4_*.npz/.json.
Only the reference's RESIDUALS of these edge kinds are pinned (its Jacobians are not a model: oracle/ba_oracle.c header);
the depth run also records what its LM loop does with them.  TEST INFRASTRUCTURE.  Usage (under gpurun):
    python oracle/run_ref_goldens_f4.py gpurun_out/golden_f4
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-bundle-adjustment_b200"))
from cugo_b200 import graphio, synth  # noqa: E402

DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver_f4")
ICP_KW = dict(n_poses=5, n_fixed=1, n_planes=400, n_lines=300, rot=0.2, trans=0.5, seed=17)   # tests regenerate the same problem


def main():
    out = sys.argv[1]
    os.makedirs(out, exist_ok=True)
    # depth: the "small" graph, its stereo edges read as (u, v, 1/z)
    g = synth.make_graph("small")
    bag = os.path.join(out, "small.bag")
    graphio.write_graph(g, bag)
    pre = os.path.join(out, "raw_depth")
    r = subprocess.run([DRIVER, "depth", bag, pre, "5"], capture_output=True, text=True, timeout=600)
    open(pre + ".log", "w").write(r.stdout + "\n--- stderr ---\n" + r.stderr)
    if r.returncode != 0:
        print("depth run failed", r.returncode, r.stderr[-2000:])
        return 1
    meta = json.load(open(pre + ".json"))
    n = g.stereo_p.shape[0]
    e2pl = np.fromfile(pre + ".depth_edge2pl.bin", np.int32).reshape(-1, 2)
    err = np.fromfile(pre + ".depth_errors.bin", np.float64).reshape(-1, 3)
    assert e2pl.shape[0] == err.shape[0] <= n
    np.savez_compressed(os.path.join(out, "ref_f4_depth_small.npz"), edge2pl=e2pl, errors=err)
    meta["trials_log"] = [ln for ln in r.stdout.splitlines() if "levenberg" in ln.lower()][:12]
    json.dump(meta, open(os.path.join(out, "ref_f4_depth_small.json"), "w"), indent=1)
    print("depth:", meta["F0"], meta["chi2"], meta["trials_log"][:3])
    # icp
    prob = synth.make_icp_problem(**ICP_KW)
    path = os.path.join(out, "p.icp")
    graphio.write_icp(prob, path)
    pre = os.path.join(out, "raw_icp")
    r = subprocess.run([DRIVER, "icp", path, pre], capture_output=True, text=True, timeout=600)
    open(pre + ".log", "w").write(r.stdout + "\n--- stderr ---\n" + r.stderr)
    if r.returncode != 0:
        print("icp run failed", r.returncode, r.stderr[-2000:])
        return 1
    meta = json.load(open(pre + ".json"))
    arr = {}
    for k, c in (("plane", 7), ("line", 10)):
        arr[k + "_meas"] = np.fromfile(pre + ".%s_meas.bin" % k, np.float64).reshape(-1, c)
        arr[k + "_errors"] = np.fromfile(pre + ".%s_errors.bin" % k, np.float64)
        arr[k + "_pose"] = np.fromfile(pre + ".%s_edge2pl.bin" % k, np.int32).reshape(-1, 2)[:, 0].copy()
    np.savez_compressed(os.path.join(out, "ref_f4_icp.npz"), **arr)
    meta["icp_kw"] = ICP_KW
    json.dump(meta, open(os.path.join(out, "ref_f4_icp.json"), "w"), indent=1)
    print("icp:", meta)
    return 0


if __name__ == "__main__":
    sys.exit(main())
