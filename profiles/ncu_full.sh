#!/bin/bash
# One `ncu --set full` capture of the kernels matching a regex (one GPU; run under gpurun, after the same command exited 0 without ncu).
#   bash profiles/ncu_full.sh <tag> <kernel regex> <skip> <count> [extra bench.py flags]
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
ncu --set full --clock-control none --import-source on -k regex:"$rx" --launch-skip $skip --launch-count $cnt -f -o gpurun_out/${tag} \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-side-shapes "$@" > gpurun_out/${tag}_ncu.log 2>&1
ls -la gpurun_out/${tag}.ncu-rep
