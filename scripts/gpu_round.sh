#!/bin/bash
# Round-end style GPU session: parity tests, both bench arms, ncu launch list + full capture.
# Usage (on the GPU box, from the repo root): bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python -m pytest tests -q -m gpu -x > $OUT/pytest_gpu.log 2>&1; echo "pytest rc $?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
python bench.py --impl reference --steps 5 --warmup 2 > $OUT/bench_reference.json 2> $OUT/bench_reference.err
python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
cat $OUT/bench_reference.json | cut -c1-600; echo; cat $OUT/bench.json | cut -c1-1500; echo
CMD="python bench.py --steps 1 --warmup 3 --cpu-iters 1"
$CMD > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 900 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
$CMD > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_schur_rows|k_schur_inv|k_build|k_chol_panel3|k_chol_update3|k_chol_update_small|k_chol_bwd3|k_chol_zinv|k_backsub|k_errors|k_update_poses" -c 80 -o $OUT/prof $CMD > $OUT/ncu_full.log 2>&1
ls -la $OUT
ncu -i $OUT/prof.ncu-rep --page raw --csv > $OUT/ncu_full_raw.csv 2> $OUT/ncu_import.err
python scripts/ncu_summary.py $OUT/ncu_full_raw.csv $OUT/ncu_full_summary.csv $OUT/ncu_traffic.json
scripts/micro/fp64_lat > $OUT/fp64_microbench.log 2>&1
bash scripts/ab_lib.sh kitti00 cuda-bundle-adjustment_b200/lib/libcugo_b200_base.so - > $OUT/panel_ab.log 2>&1
python scripts/ab_opts.py kitti00 - panel_dc=0 panel_rb=10 upd_small=0 schur_mma=1 pdl=0 > $OUT/ab_options.log 2>&1
python scripts/gpu_profile.py venice > $OUT/venice_per_kernel.log 2>&1
python scripts/bench_poseonly.py > $OUT/poseonly_bench.log 2>&1      # row f4: ICP scan / window, motion-only KITTI-00 shape
# gpurun brings back at most 64 MiB: the report itself stays on the box, its CSV exports travel
rm -f $OUT/prof.ncu-rep; gzip -f $OUT/ncu_full_raw.csv
