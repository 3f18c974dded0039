#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_search.py tests/test_gpu_e2e.py tests/test_compat.py tests/test_gpu_net.py tests/test_gpu_selfplay.py tests/test_facade.py -x -q -m gpu > gpurun_out/tree_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/tree_pytest.log
for T in 1024 16384; do timeout 300 python tools/tree_profile.py c4d $T 50 3 | tail -1; done
CMD="python tools/tree_profile.py c4d 16384 50 2"
timeout 600 ncu --metrics gpu__time_duration.sum,sm__inst_executed.sum --clock-control none -k regex:tree_step --launch-skip 75 -c 20 --csv --log-file gpurun_out/tree_inst.csv $CMD > /dev/null 2>&1
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/tree_inst.csv') if not l.startswith('==')]
d={}
for r in csv.DictReader(lines):
    d.setdefault(r['Metric Name'],[]).append(float(r['Metric Value'].replace(',','')))
for k,v in d.items(): print(k, sum(v)/len(v))
PY
for g in ttt c4r; do echo -n "$g search: "; timeout 300 python tools/root_timing.py $g 2>&1 | grep "noise 0" | tail -1 | cut -c1-120; done
