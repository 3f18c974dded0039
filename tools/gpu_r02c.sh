#!/bin/bash
# kernel iteration: bf16 parity + tower timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_net.py tests/test_gpu_e2e.py tests/test_gpu_pipeline.py -x -q -k "bf16 or replays or pipeline or trained" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/quick_pytest.log
for g in c4d c4r ttt ale; do timeout 300 python tools/profile_recurrent.py $g 1024 50 2>&1 | tail -2; done
MZB_FUSED_NO_STATIC_MASK=1 timeout 300 python tools/profile_recurrent.py c4d 1024 50 2>&1 | tail -1
MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 1024 2 2>&1 | grep "fused" | tail -34
