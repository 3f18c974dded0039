#!/bin/bash
# 16x256b read-out variant (library tools/bin/libmzb200_16x256.so) against the tree's library: parity tests, pass times, phase clocks
MZB200_LIB=$PWD/tools/bin/libmzb200_16x256.so timeout 900 python -m pytest tests/test_gpu_net.py tests/test_gpu_pipeline.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -3
for rep in 1 2; do
for spec in "c4d 592" "c4d 1024" "ale 1024" "c4r 1024" "ttt 1024"; do set -- $spec
  MZB200_LIB=$PWD/tools/bin/libmzb200_16x256.so timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/16x256   $1 T=$2 /"
  timeout 300 python tools/profile_recurrent.py $1 $2 300 2>&1 | grep recurrent | sed "s/^/base     $1 T=$2 /"
done
done
MZB200_LIB=$PWD/tools/bin/libmzb200_16x256.so MZB_FUSED_TIMING=1 timeout 300 python tools/profile_recurrent.py c4d 592 2 2>&1 | grep "fused" | tail -24 | head -9
for cfg in c4d ale c4r; do
MZB200_LIB=$PWD/tools/bin/libmzb200_16x256.so timeout 300 python tools/dbg_split_diff.py $cfg /tmp/v_$cfg.npz; timeout 300 python tools/dbg_split_diff.py $cfg /tmp/b_$cfg.npz
python -c "
import numpy as np
a=np.load('/tmp/b_$cfg.npz'); b=np.load('/tmp/v_$cfg.npz')
print('$cfg bit-identical:', {k: bool(np.array_equal(a[k], b[k])) for k in a.files})"
done
