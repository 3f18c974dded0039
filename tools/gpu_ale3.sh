#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py -x -q > gpurun_out/pipe_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pipe_pytest.log
echo -n "ale alternate:  "; timeout 300 python tools/root_timing.py ale 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
echo -n "ale in phase:   "; MZB_PIPELINE_ALTERNATE=0 timeout 300 python tools/root_timing.py ale 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
echo -n "ale alt 3 parts:"; MZB_PIPELINE_PARTS=3 timeout 300 python tools/root_timing.py ale 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
echo -n "c4d default:    "; timeout 300 python tools/root_timing.py c4d 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
echo -n "c4d alternate:  "; MZB_PIPELINE_ALTERNATE=1 timeout 300 python tools/root_timing.py c4d 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
echo -n "c4r alternate:  "; MZB_PIPELINE_ALTERNATE=1 timeout 300 python tools/root_timing.py c4r 2>&1 | grep "noise 0" | tail -1 | cut -c1-130
MZB_PIPELINE_TRACE=1 timeout 300 python tools/root_timing.py ale 2>&1 | grep -E "pipeline\] sim  [3-5]" | tail -3
