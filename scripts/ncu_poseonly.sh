#!/bin/bash
# ncu capture of the row-f4 kernels (one B200): launch list + full sets of the ICP kernels and the block-diagonal solve.
# Usage (GPU box, repo root): bash scripts/ncu_poseonly.sh [tag]
TAG=${1:-r05}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python scripts/bench_poseonly.py icp_window"
$CMD > $OUT/poseonly_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_icp|k_po_solve|k_lm|k_update_poses|k_errors" -c 300 --csv --log-file $OUT/poseonly_launches.csv $CMD > $OUT/poseonly_ncu_launches.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_icp_build|k_icp_errors|k_icp_add|k_po_solve" -s 8 -c 8 -o $OUT/poseonly_prof $CMD > $OUT/poseonly_ncu_full.log 2>&1
ncu -i $OUT/poseonly_prof.ncu-rep --page raw --csv > $OUT/poseonly_ncu_full_raw.csv 2> $OUT/poseonly_ncu_import.err
python scripts/ncu_summary.py $OUT/poseonly_ncu_full_raw.csv $OUT/poseonly_ncu_full_summary.csv $OUT/poseonly_ncu_traffic.json
rm -f $OUT/poseonly_prof.ncu-rep; gzip -f $OUT/poseonly_ncu_full_raw.csv
ls -la $OUT | head -20
cat $OUT/poseonly_ncu_traffic.json | head -40
