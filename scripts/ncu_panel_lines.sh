#!/bin/bash
# one ncu --set full capture of a panel launch (a 16-wide level) + per-source-line stall summary
TAG=${1:-x}; RX=${2:-k_chol_panel3}; SKIP=${3:-36}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python scripts/gpu_profile.py kitti00"
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c 1 -o $OUT/prof $CMD > $OUT/ncu.log 2>&1
python scripts/ncu_lines.py $OUT/prof.ncu-rep 60 > $OUT/lines.txt 2>&1
ncu -i $OUT/prof.ncu-rep --page raw --csv > $OUT/raw.csv 2> $OUT/import.err
gzip -f $OUT/raw.csv
rm -f $OUT/prof.ncu-rep
head -70 $OUT/lines.txt
