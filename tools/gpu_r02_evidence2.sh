#!/bin/bash
# Round-2 evidence, part 2: --set full captures of the downsample-head kernels and the head MLP kernel; refreshed bf16 error table
mkdir -p gpurun_out
CMD="python tools/profile_recurrent.py ale 1024 2"
$CMD > gpurun_out/plain_ale.log 2>&1 || { echo "plain failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_first -c 1 -f -o gpurun_out/r02_conv_first $CMD > /dev/null 2>&1; echo "cf rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_band2 -c 11 -f -o gpurun_out/r02_conv_band2 $CMD > /dev/null 2>&1; echo "cb rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_heads -s 2 -c 1 -f -o gpurun_out/r02_mlp_heads $CMD > /dev/null 2>&1; echo "mlp rc=$?"
timeout 900 python tools/bf16_error.py --out gpurun_out/bf16_error_r02.json > gpurun_out/bf16_error.log 2>&1; echo "bf16_error rc=$?"; tail -3 gpurun_out/bf16_error.log
ls -la gpurun_out/*.ncu-rep | tail -4
