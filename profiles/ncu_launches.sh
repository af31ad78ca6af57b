#!/bin/bash
# Per-launch duration and DRAM bytes of every kernel of a short bench run (one GPU; run under gpurun).
#   bash profiles/ncu_launches.sh <tag> [extra bench.py flags]
# The program is first run without ncu (must exit 0), then under ncu; the CSV lands in gpurun_out/<tag>_launches.csv.
tag=$1; shift
set -e
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-side-shapes "$@" > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-side-shapes "$@" > gpurun_out/${tag}_ncu.log 2>&1
