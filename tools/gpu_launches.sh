#!/bin/bash
# ncu launch list only.  Args: game trees sims
mkdir -p gpurun_out
CMD="python tools/profile_recurrent.py ${1:-c4d} ${2:-1024} ${3:-4}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"; tail -2 gpurun_out/plain.log
