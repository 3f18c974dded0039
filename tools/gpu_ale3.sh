#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_net.py tests/test_gpu_e2e.py -x -q > gpurun_out/pipe_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pipe_pytest.log
for S in 1 2 4 8; do echo -n "ale splits $S: "; MZB_MLP_SPLITS=$S timeout 300 python tools/profile_recurrent.py ale 1024 50 2>&1 | grep recurrent; done
echo -n "ale search: "; timeout 300 python tools/root_timing.py ale 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
CMD="python tools/profile_recurrent.py ale 1024 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mlp_heads -c 8 --csv --log-file gpurun_out/mlp.csv $CMD > /dev/null 2>&1; grep -c mlp_heads gpurun_out/mlp.csv; grep mlp_heads gpurun_out/mlp.csv | tail -3 | cut -d, -f 5,12-
