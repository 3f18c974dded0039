#!/bin/bash
# Round-2 pipeline experiments (results: profiles/r02_pipeline_experiments.txt).  Needs tools/bin/libmzb200_r152.so =
# the library built with -DMZB_TOWER_MAXNREG=152 (see tower_fused.cuh) and tools/bin/coreside_probe (tools/coreside_probe.cu).
L="MZB200_LIB=$PWD/tools/bin/libmzb200_r152.so"
tools/bin/coreside_probe
bash tools/gpu_tile_time.sh
for R in 168 152; do
  if [ $R = 168 ]; then E=""; else E="$L"; fi
  echo "== tower regs $R, equal cut";    env $E timeout 300 python tools/root_timing.py 2>&1 | grep "noise 0" | tail -1
  echo "== tower regs $R, SM-aware cut"; env $E MZB_PIPELINE_SM_CUT=1 timeout 300 python tools/root_timing.py 2>&1 | grep "noise 0" | tail -1
  for K in 3 4 6 8; do echo -n "regs $R parts $K: "; env $E MZB_PIPELINE_PARTS=$K timeout 300 python tools/root_timing.py 2>&1 | grep "noise 0" | tail -1 | cut -c1-110; done
done
env $L MZB_PIPELINE_SM_CUT=1 MZB_TILE_TRACE=1 timeout 300 python tools/root_timing.py 2>&1 | grep -E "tiles\] launch  (4[0-7])" | tail -8
MZB_TILE_TRACE=1 timeout 300 python tools/root_timing.py 2>&1 | grep -E "tiles\] launch  (4[0-7])" | tail -8
