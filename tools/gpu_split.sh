#!/bin/bash
# staggered chunk order of the two warps of a lane quarter (dbg flag 32 = lock step as before)
for fl in 32 0; do
for spec in "c4d 592" "c4d 1024" "ale 1024" "c4r 1024" "ttt 1024"; do set -- $spec
  MZB_FUSED_DBGFLAGS=$fl timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/flags=$fl $1 T=$2 /"
done
done
for fl in 32 0; do
MZB_FUSED_DBGFLAGS=$fl MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -9 | sed "s/^/flags=$fl /"
done
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_gpu_pipeline.py -x -q 2>&1 | tail -2
