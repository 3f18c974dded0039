#!/bin/bash
# ALE-shaped initial inference: parity tests, then the launch list of one search
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_net.py tests/test_gpu_e2e.py -x -q > gpurun_out/ale_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/ale_pytest.log
CMD="python tools/profile_recurrent.py ale 1024 2"
$CMD 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02b_launches_ale.csv $CMD > gpurun_out/ncu_launches_ale.log 2>&1
echo "ale launch list rc=$?"
