#!/bin/bash
# Round-2 ncu evidence: --set full of tower_fused_kernel for c4r / ale / ttt at 1024 states and c4d at 4096,
# --set full of tree_step_kernel at 1024 and 16384 trees, launch list of an ALE-shaped search.  Each capture follows a plain run.
mkdir -p gpurun_out
for spec in "c4r 1024" "ale 1024" "ttt 1024" "c4d 4096"; do
  set -- $spec
  CMD="python tools/profile_recurrent.py $1 $2 4"
  $CMD > gpurun_out/plain_$1_$2.log 2>&1 || { echo "plain $1 $2 failed"; tail -3 gpurun_out/plain_$1_$2.log; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:tower_fused -s 3 -c 1 -f -o gpurun_out/r02_tower_$1_T$2 $CMD > gpurun_out/ncu_$1_$2.log 2>&1
  echo "tower $1 $2 rc=$?"; tail -1 gpurun_out/plain_$1_$2.log
done
for T in 1024 16384; do
  CMD="python tools/tree_profile.py c4d $T 50 2"
  $CMD > gpurun_out/tree_plain_$T.log 2>&1 || { echo "plain tree $T failed"; continue; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:tree_step -s 75 -c 1 -f -o gpurun_out/r02_tree_step_T$T $CMD > gpurun_out/ncu_tree_$T.log 2>&1
  echo "tree $T rc=$?"; tail -1 gpurun_out/tree_plain_$T.log
done
CMD="python tools/profile_recurrent.py ale 1024 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_ale.csv $CMD > gpurun_out/ncu_launches_ale.log 2>&1
echo "ale launch list rc=$?"
ls -la gpurun_out/*.ncu-rep
