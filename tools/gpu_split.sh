#!/bin/bash
# staggered half tiles: bit-identity against the unsplit kernel, then the bench loop by (head, tail) slots per half
mkdir -p gpurun_out
for cfg in c4d ale; do
  MZB_TOWER_SPLIT=0,0 timeout 300 python tools/dbg_split_diff.py $cfg /tmp/a_$cfg.npz
  for sp in 1,1 2,2 3,3; do
    MZB_TOWER_SPLIT=$sp timeout 300 python tools/dbg_split_diff.py $cfg /tmp/b_$cfg.npz
    python -c "
import numpy as np,sys
a=np.load('/tmp/a_$cfg.npz'); b=np.load('/tmp/b_$cfg.npz')
print('$cfg split=$sp', {k: bool(np.array_equal(a[k], b[k], equal_nan=False)) for k in a.files})"
  done
done
for sp in 0,0 1,1 2,2 1,2 2,1 0,1 0,2; do
  MZB_TOWER_SPLIT=$sp timeout 300 python bench.py --steps 60 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('split=$sp c4d bench', round(d['ms_per_step'],3), 'ms/step', round(d['value']/1e6,3), 'M sims/s', d['clocks']['sm_mhz'])"
done
