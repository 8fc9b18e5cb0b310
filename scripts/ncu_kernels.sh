#!/bin/bash
# ncu --set full capture of selected trial kernels (one launch each) + raw CSV export.
# usage (under gpurun): bash scripts/ncu_kernels.sh TAG "regex of kernel names" [count]
TAG=$1; RX=$2; CNT=${3:-6}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python scripts/gpu_profile.py kitti00"
$CMD > $OUT/plain.log 2>&1 || { tail -5 $OUT/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 4 -c $CNT -o $OUT/prof $CMD > $OUT/ncu.log 2>&1
ncu -i $OUT/prof.ncu-rep --page raw --csv > $OUT/raw.csv 2> $OUT/import.err
ncu -i $OUT/prof.ncu-rep --page details --csv > $OUT/details.csv 2>> $OUT/import.err
rm -f $OUT/prof.ncu-rep
gzip -f $OUT/raw.csv $OUT/details.csv
ls -la $OUT
