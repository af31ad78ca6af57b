#!/bin/bash
# One `ncu --set full` capture of the kernels matching a regex (one GPU; run under gpurun, after the same command exited 0 without ncu).
#   bash profiles/ncu_full.sh <tag> <kernel regex> <skip> <count> [extra bench.py flags]
# Leaves gpurun_out/<tag>_details.txt, <tag>_raw.csv and (KEEP_REP=1) the .ncu-rep itself (gpurun copies back at most 64 MiB).
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
ncu --set full --clock-control none --import-source on -k regex:"$rx" --launch-skip $skip --launch-count $cnt -f -o /tmp/${tag} \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-side-shapes "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu -i /tmp/${tag}.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
ncu -i /tmp/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
if [ "$KEEP_REP" = "1" ]; then cp /tmp/${tag}.ncu-rep gpurun_out/; fi
ls -la gpurun_out/${tag}_details.txt
