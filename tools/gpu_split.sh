#!/bin/bash
# experiment: X with 128-byte aligned planes (R multiple of 8): what do the epilogue's bank conflicts cost?
for rep in 1 2; do
for al in 0 1; do
  if [ $al = 1 ]; then export MZB_X_ALIGN8=1; else unset MZB_X_ALIGN8; fi
  for spec in "c4d 592" "c4d 1024"; do set -- $spec
    timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep -E "recurrent|rror" | sed "s/^/align8=$al $1 T=$2 /"
  done
done
done
export MZB_X_ALIGN8=1
MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -6
timeout 600 python -m pytest tests/test_gpu_net.py -x -q -k "c4d or c4r" 2>&1 | tail -2
