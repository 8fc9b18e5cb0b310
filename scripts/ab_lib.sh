#!/bin/bash
# A/B of two builds of the library on one box: scripts/ab_lib.sh <workload> <other .so> [variants...]
wl=${1:-kitti00}; other=$2; shift 2
echo "== current build"; python scripts/ab_opts.py $wl "$@"
echo "== $other"; CUGO_B200_LIB=$other python scripts/ab_opts.py $wl "$@"
echo "== current build again"; python scripts/ab_opts.py $wl "$@"
