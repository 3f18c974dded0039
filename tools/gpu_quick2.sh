#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_net.py tests/test_gpu_e2e.py tests/test_gpu_pipeline.py tests/test_gpu_search.py -x -q > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/quick_pytest.log
for g in c4r ttt c4d ale; do echo -n "$g: "; timeout 300 python tools/profile_recurrent.py $g 1024 50 2>&1 | grep recurrent; echo -n "$g no dup: "; MZB_FUSED_NO_DUP=1 timeout 300 python tools/profile_recurrent.py $g 1024 50 2>&1 | grep recurrent; done
for g in c4r ttt; do echo -n "$g search: "; timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c1-120; done
