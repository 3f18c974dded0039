#!/bin/bash
# round 2, session d: whole GPU suite, then the SM-aware pipeline cut against the equal cut, root-phase timing
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02d_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r02d_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_c4d.json 2> gpurun_out/r02d_bench_c4d.err; echo "bench rc=$?"
cat gpurun_out/r02d_bench_c4d.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','e2e','roofline')}); print(d.get('workloads'))"
MZB_PIPELINE_EQUAL_CUT=1 MZB_PIPELINE_FULL_TILES=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('equal cut', {k:d[k] for k in ('value','ms_per_step')})"
timeout 300 python tools/root_timing.py 2>&1 | tail -4
