#!/bin/bash
# `ncu --set full` of kernels matching a regex in an arbitrary command:  bash profiles/ncu_any.sh <tag> <regex> <skip> <count> <command...>
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
ncu --set full --clock-control none --import-source on -k regex:"$rx" --launch-skip $skip --launch-count $cnt -f -o /tmp/${tag} "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu -i /tmp/${tag}.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
ncu -i /tmp/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
if [ "$KEEP_REP" = "1" ]; then cp /tmp/${tag}.ncu-rep gpurun_out/; fi
ls -la gpurun_out/${tag}_details.txt
