#!/bin/bash
# ncu launch list + one full capture of the dominant kernel.  Args: kernel regex (default umma_conv3)
K=${1:-umma_conv3}
mkdir -p gpurun_out
CMD="python tools/profile_recurrent.py c4d 1024 4"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:$K -s 40 -c 3 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
cat gpurun_out/plain.log | tail -5
