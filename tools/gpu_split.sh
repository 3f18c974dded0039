#!/bin/bash
# three warpgroups + setmaxnreg (service 72 / epilogue 216 registers): parity tests, pass times
timeout 900 python -m pytest tests/test_gpu_net.py tests/test_gpu_pipeline.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -3
for rep in 1 2; do
for spec in "c4d 592" "c4d 1024" "ale 1024" "c4r 1024" "ttt 1024"; do set -- $spec
  timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/dblbuf   $1 T=$2 /"
  MZB200_LIB=$PWD/tools/bin/libmzb200_base.so timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/base     $1 T=$2 /"
done
done
MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -9
