#!/bin/bash
# production library + the diagnostic twin (phase clocks), from any directory
set -e
P=/root/repo/cuda-bundle-adjustment_b200
make -C $P 2>&1 | grep -E "error|Error" && exit 1
nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -DCUGO_B200_DIAG -shared -o $P/lib/libcugo_b200_diag.so $P/csrc/engine.cu 2>&1 | grep -E "error" && exit 1
ls -la $P/lib/libcugo_b200.so $P/lib/libcugo_b200_diag.so
grep -B1 -A1 "k_chol_panel3ILb1" $P/lib/ptxas.log | grep -E "Used|spill"
