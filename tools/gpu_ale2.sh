#!/bin/bash
# ncu --set full of the downsample head's kernels (one ALE-shaped search: 1 conv_first launch, 11 band launches)
mkdir -p gpurun_out
CMD="python tools/profile_recurrent.py ale 1024 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_first -c 1 -f -o gpurun_out/r02_conv_first $CMD > gpurun_out/ncu_cf.log 2>&1; echo "cf rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_band -c 11 -f -o gpurun_out/r02_conv_band $CMD > gpurun_out/ncu_cb.log 2>&1; echo "cb rc=$?"
ls -la gpurun_out/*.ncu-rep | tail -3
