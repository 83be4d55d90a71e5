#!/bin/bash
# ncu launch list (per-launch gpu__time_duration, no clock control) of a command, written as CSV under gpurun_out/.
#   scripts/ncu_launches.sh <tag> <command...>      run on the GPU box (gpurun); never a multi-rank command
tag=$1; shift
ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_COUNT:-400} --csv --log-file gpurun_out/${tag}_launches.csv "$@" > gpurun_out/${tag}_ncu.log 2>&1
python scripts/ncu_launch_summary.py gpurun_out/${tag}_launches.csv > gpurun_out/${tag}_launches_summary.json
cat gpurun_out/${tag}_launches_summary.json
