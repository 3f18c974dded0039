#!/bin/bash
# ncu evidence for one round.  Args: kernel regex for the full capture (default tower_fused).
#   1. plain bench run (must exit 0 first)           -> gpurun_out/prof_bench.json
#   2. ncu launch list of the SAME bench command     -> gpurun_out/launches_bench.csv
#   3. ncu --set full of the dominant kernel         -> gpurun_out/prof.ncu-rep (short driver, fewer launches to skip)
K=${1:-tower_fused}
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$BENCH > gpurun_out/prof_bench.json 2> gpurun_out/prof_bench.err || { echo "plain bench failed"; tail -5 gpurun_out/prof_bench.err; exit 1; }
cat gpurun_out/prof_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_bench.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
CMD="python tools/profile_recurrent.py c4d 1024 4"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 3 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
tail -5 gpurun_out/plain.log
