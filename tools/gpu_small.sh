#!/bin/bash
# where a simulation's time goes for the small towers (ttt, c4r): pipeline timeline + parts sweep
for G in ttt c4r; do
  echo "== $G"; MZB_PIPELINE_TRACE=1 timeout 300 python tools/root_timing.py $G 2>&1 | grep -E "pipeline\] (parts|sim  [2-4])|noise 0" | tail -5
  for K in 1 3 4; do echo -n "$G parts $K: "; if [ $K = 1 ]; then E="MZB_PIPELINE_MIN_TREES=0"; else E="MZB_PIPELINE_PARTS=$K"; fi; env $E timeout 300 python tools/root_timing.py $G 2>&1 | grep "noise 0" | tail -1 | cut -c1-120; done
done
