#!/bin/bash
# quick GPU check of the bf16 path: parity tests that touch it + a recurrent timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_gpu_e2e.py -x -q -k "bf16 or replays" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/quick_pytest.log
for g in c4d c4r ttt; do timeout 300 python tools/profile_recurrent.py $g 1024 50 2>&1 | tail -3; done
timeout 300 python tools/profile_recurrent.py c4d 4096 50 2>&1 | tail -3
