#!/bin/bash
# band kernel v2 (PKM): parity, search timing, then where a band's time goes — launch lists with parts of the work switched off
# (MZB_BAND_DBG bits: 1 weight stages shrink to 256 B, 2 no epilogue work, 4 no band loads; results are wrong in those runs)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_gpu_pipeline.py -x -q -k "ale or tinyds or ragged" > gpurun_out/band_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/band_pytest.log
echo -n "search: "; timeout 300 python tools/root_timing.py ale 2>&1 | grep "noise 0" | tail -1 | cut -c1-110
CMD="python tools/profile_recurrent.py ale 1024 2"
for D in ${BAND_DBG_LIST:-0 1 2 4 6}; do
  MZB_BAND_DBG=$D timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"conv_band|conv_first|avgpool" -c 14 --csv --log-file gpurun_out/band2_$D.csv $CMD > /dev/null 2>&1
  echo -n "dbg $D (us): "; grep -E "conv_band|conv_first|avgpool" gpurun_out/band2_$D.csv | awk -F'","' '{gsub(/"/,"",$NF); printf "%d ", $NF/1000}'; echo
done
