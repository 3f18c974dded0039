#!/bin/bash
# staggered half tiles incl. duplicated-row towers (c4r): pipeline parity tests, then pass times by setting
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_net.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -3
for sp in 0,0 1,1 2,2 1,2; do
  for spec in "c4r 592" "c4d 592" "ale 592"; do set -- $spec
    MZB_TOWER_SPLIT=$sp timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/split=$sp $1 T=$2 /"
  done
done
MZB_TOWER_SPLIT=1,1 MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4r 592 2 2>&1 | grep "fused" | tail -24 | head -12
