#!/bin/bash
# whole GPU suite + ALE launch list + default bench
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02f_pytest.log
CMD="python tools/profile_recurrent.py ale 1024 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02f_launches_ale.csv $CMD > /dev/null 2>&1
python tools/summarize_launches.py gpurun_out/r02f_launches_ale.csv 2>/dev/null | head -8
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r02f_bench_c4d.json 2> gpurun_out/r02f_bench_c4d.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02f_bench_c4d.json'))
print({k:d[k] for k in ('value','ms_per_step','steps')}, d['e2e']['value'], d['roofline']['frac'], d['roofline']['launch_ms'])
for k,v in d.get('workloads',{}).items(): print(k, {x:v[x] for x in ('value','e2e','ms_per_step','tower_frac_of_burst_peak','root_ms_per_step')})
"
