#!/bin/bash
# A/B of the engine's environment switches on one box: bash scripts/ab_env.sh "VAR=0" "VAR2=0" ...
for cfg in "" "$@"; do
  echo "== ${cfg:-default}"
  env $cfg python scripts/gpu_profile.py ${SHAPE:-kitti00} 2>&1 | grep "optimize(10)" | tail -3
done
