#!/bin/bash
# panel-loop work: parity of the current build, A/B against lib/libcugo_b200_base.so, DIAG trace of the loop
tag=${1:-x}
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or switches or oracle" 2>&1 | tail -2
bash scripts/ab_lib.sh kitti00 cuda-bundle-adjustment_b200/lib/libcugo_b200_base.so - 2>&1 | grep -v "^$"
CUGO_B200_DEBUG=1 CUGO_B200_LIB=cuda-bundle-adjustment_b200/lib/libcugo_b200_diag.so python scripts/gpu_profile.py kitti00 > gpurun_out/${tag}_diag.log 2>&1
grep "decoupled panel loop" -A17 gpurun_out/${tag}_diag.log | cut -c1-150
