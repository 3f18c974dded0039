#!/bin/bash
# whole GPU suite, bench lines (default run incl. workloads), root timing with 1 / 4 noise threads
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02e_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02e_pytest.log
MZB_NOISE_THREADS=1 timeout 300 python tools/root_timing.py 2>&1 | grep "noise 1" | tail -1 | cut -c1-140
timeout 300 python tools/root_timing.py 2>&1 | grep "noise 1" | tail -1 | cut -c1-140
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/r02e_bench_c4d.json 2> gpurun_out/r02e_bench_c4d.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02e_bench_c4d.json'))
print({k:d[k] for k in ('value','ms_per_step','steps')}, d['e2e']['value'], d['roofline']['frac'], d['roofline']['launch_ms'])
print(d.get('sustained'))
for k,v in d.get('workloads',{}).items(): print(k, {x:v[x] for x in ('value','e2e','ms_per_step','tower_frac_of_burst_peak','root_ms_per_step','traffic')})
"
