#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/graph_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/graph_pytest.log
for g in c4d c4r ttt ale; do
  echo -n "$g graph:    "; timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c1-125
  echo -n "$g no graph: "; MZB_NO_GRAPH=1 timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c1-125
done
