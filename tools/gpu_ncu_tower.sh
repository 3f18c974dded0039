#!/bin/bash
# ncu --set full of tower_fused_kernel with source-level stall samples (one plain run first, as the profiling recipe asks)
G=${1:-c4d}
mkdir -p gpurun_out
CMD="python tools/profile_recurrent.py $G 1024 4"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tower_fused -s 3 -c 2 -f -o gpurun_out/prof_$G $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
tail -3 gpurun_out/plain.log; tail -3 gpurun_out/ncu_full.log
