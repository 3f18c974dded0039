#!/bin/bash
# One GPU-box visit: parity tests, bench, launch list.  Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 python bench.py --steps 3 --warmup 3 --workload c4r --no-cpu-baseline > gpurun_out/bench_c4r.json 2> gpurun_out/bench_c4r.err; cat gpurun_out/bench_c4r.json
for w in reanalyze ttt ale; do timeout 600 python bench.py --steps 2 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; cut -c1-400 gpurun_out/bench_$w.json; tail -2 gpurun_out/bench_$w.err; done
