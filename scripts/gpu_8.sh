#!/bin/bash
# 8-GPU session: city shape with the factorisation distributed by subtree (and replicated, for comparison), then the
# driver's weak-scaling line of the KITTI-00 shape.  usage (gpurun --gpus 8): bash scripts/gpu_8.sh TAG
TAG=${1:-r02}
O=gpurun_out
CUGO_B200_DEBUG=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 \
  bench.py --gpus 8 --workload city --scaling strong --skip-other --dist-compare --steps 3 --warmup 3 > $O/${TAG}_city_8gpu.json 2> $O/${TAG}_city_8gpu.err
grep -E "distributed factorisation" $O/${TAG}_city_8gpu.err | sort | uniq | cut -c1-300 | head -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 \
  bench.py --gpus 8 --steps 10 --warmup 3 > $O/${TAG}_kitti00_8gpu.json 2> $O/${TAG}_kitti00_8gpu.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29516 \
  bench.py --gpus 8 --workload venice --scaling strong --skip-other --steps 5 --warmup 3 > $O/${TAG}_venice_8gpu.json 2> $O/${TAG}_venice_8gpu.err
python - <<PY
import json
for n in ("city", "kitti00", "venice"):
    try:
        d = json.load(open("$O/${TAG}_%s_8gpu.json" % n))
        print(n, {k: d.get(k) for k in ("value", "ms_per_step", "scaling", "parity_vs_1gpu", "dist_compare", "other_scaling")})
    except Exception as e:
        print(n, "failed", e)
PY
