#!/bin/bash
# epilogue column range trimmed behind a game's last real pixel: parity tests, pass times against the previous library
timeout 900 python -m pytest tests/test_gpu_net.py tests/test_gpu_pipeline.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -3
for rep in 1 2; do
for spec in "ale 592" "ale 1024" "c4d 592" "ttt 1024"; do set -- $spec
  timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/trim     $1 T=$2 /"
  MZB200_LIB=$PWD/tools/bin/libmzb200_base.so timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/base     $1 T=$2 /"
done
done
timeout 300 python tools/root_timing.py ale 1024 2>&1 | tail -3
