#!/bin/bash
# tree steps on high-priority streams (MZB_TREE_PRIORITY=1) with the loop as a graph, by workload
for g in c4d c4r ttt ale; do
  echo -n "$g baseline:        "; timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c30-100
  for K in 2 4 6; do echo -n "$g prio parts $K:    "; MZB_TREE_PRIORITY=1 MZB_PIPELINE_PARTS=$K timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c30-100; done
done
MZB_TREE_PRIORITY=1 MZB_PIPELINE_PARTS=4 timeout 600 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -2
