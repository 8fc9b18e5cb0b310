#!/bin/bash
# 2-GPU session: the 2-GPU tests, the driver's weak-scaling line of the KITTI-00 shape (with the strong line in other_scaling) and the
# city shape with the factorisation by subtree against replicated.  usage (gpurun --gpus 2): bash scripts/gpu_2.sh TAG
TAG=${1:-r02}
O=gpurun_out
python -m pytest tests -x -q -m gpu -k "two_gpus" 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29524 \
  bench.py --gpus 2 --steps 10 --warmup 3 > $O/${TAG}_kitti00_2gpu.json 2> $O/${TAG}_kitti00_2gpu.err
CUGO_B200_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29525 \
  bench.py --gpus 2 --workload city --scaling strong --skip-other --dist-compare --steps 3 --warmup 3 > $O/${TAG}_city_2gpu.json 2> $O/${TAG}_city_2gpu.err
grep -E "distributed factorisation" $O/${TAG}_city_2gpu.err | sort | uniq | cut -c1-300 | head -2
python - <<PY
import json
for n in ("kitti00", "city"):
    try:
        d = json.load(open("$O/${TAG}_%s_2gpu.json" % n))
        print(n, {k: d.get(k) for k in ("value", "ms_per_step", "scaling", "parity_vs_1gpu", "dist_compare", "other_scaling")})
    except Exception as e:
        print(n, "failed", e)
PY
