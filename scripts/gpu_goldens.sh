#!/bin/bash
# Regenerates tests/golden/ from the unmodified reference on the GPU box, then runs the tests that use them.
# usage (under gpurun): bash scripts/gpu_goldens.sh [case ...]
set -x
mkdir -p gpurun_out/golden
python oracle/run_ref_goldens.py gpurun_out/golden "$@" > gpurun_out/golden/run.log 2>&1
tail -15 gpurun_out/golden/run.log
cp gpurun_out/golden/ref_*.json gpurun_out/golden/ref_*.npz tests/golden/
python -m pytest tests/test_oracle_golden_cpu.py -x -q 2>&1 | tail -8
python -m pytest tests -x -q -m gpu 2>&1 | tail -15
