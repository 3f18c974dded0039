#!/bin/bash
# Tree kernels alone under ncu: per-launch duration + DRAM bytes (achieved HBM GB/s), second search only.  Args: trees
mkdir -p gpurun_out
T=${1:-1024}
CMD="python tools/tree_profile.py c4d $T 50 2"
$CMD > gpurun_out/tree_plain_$T.log 2>&1 && cat gpurun_out/tree_plain_$T.log &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none --launch-skip 160 -c 150 --csv --log-file gpurun_out/tree_launches_$T.csv $CMD > gpurun_out/tree_ncu_$T.log 2>&1
echo "tree launch list T=$T rc=$?"
