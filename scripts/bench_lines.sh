#!/bin/bash
# The other single-GPU bench lines (one JSON line each) from one script: KITTI-07, Venice, USE_FLOAT32 storage mode, KITTI-00 + Huber.
# usage (on a B200, from the repo root): bash scripts/bench_lines.sh TAG   ->  gpurun_out/TAG_bench_<name>.json
TAG=${1:-r02}
O=gpurun_out
python bench.py --workload kitti07 --steps 10 --warmup 3 > $O/${TAG}_bench_kitti07.json 2> $O/${TAG}_bench_kitti07.err
python bench.py --workload venice --steps 5 --warmup 3 --cpu-iters 1 > $O/${TAG}_bench_venice.json 2> $O/${TAG}_bench_venice.err
python bench.py --f32 --steps 10 --warmup 3 > $O/${TAG}_bench_kitti00_f32.json 2> $O/${TAG}_bench_kitti00_f32.err
python bench.py --robust huber --delta 2.5 --steps 10 --warmup 3 > $O/${TAG}_bench_kitti00_huber.json 2> $O/${TAG}_bench_kitti00_huber.err
python - <<PY
import json
for n in ("kitti07", "venice", "kitti00_f32", "kitti00_huber"):
    try:
        d = json.loads(open("$O/${TAG}_bench_%s.json" % n).read().strip().splitlines()[-1])
        print(n, {k: d.get(k) for k in ("value", "ms_per_step", "chi2_rel_err_vs_cpu_oracle")}, d["e2e"]["ms_per_step"], d["roofline"]["kernel"], d["roofline"]["frac"])
    except Exception as e:
        print(n, "failed", e)
PY
