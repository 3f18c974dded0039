#!/bin/bash
# end-of-round evidence: GPU suite, smoke(), default bench line, launch list of the bench command
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02l_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02l_bench_c4d.json 2> gpurun_out/r02l_bench_c4d.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02l_bench_c4d.json'))
print({k:d[k] for k in ('value','ms_per_step','steps','gpu_launches')}, 'e2e', d['e2e']['value'], 'frac', d['roofline']['frac'], d['clocks'])
print('cpu', d.get('cpu_baseline'))
for k,v in d.get('workloads',{}).items(): print(k, {x:v[x] for x in ('value','e2e','ms_per_step','tower_frac_of_burst_peak','root_ms_per_step')})
"
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02l_launches_bench_c4d.csv $BENCH > /dev/null 2>&1; echo "launch list rc=$?"
python tools/summarize_launches.py gpurun_out/r02l_launches_bench_c4d.csv 2>/dev/null | head -8
