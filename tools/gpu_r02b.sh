#!/bin/bash
# Round 2, visit B: full GPU parity suite, bf16 error table on real positions, reanalyze workload
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 1200 python tools/bf16_error.py --out gpurun_out/bf16_error_r02.json > gpurun_out/bf16_error.log 2>&1; echo "bf16_error rc=$?"; tail -9 gpurun_out/bf16_error.log
timeout 600 python bench.py --steps 5 --warmup 3 --workload reanalyze --no-cpu-baseline --no-extras > gpurun_out/bench_reanalyze.json 2> gpurun_out/bench_reanalyze.err; echo "reanalyze rc=$?"; cut -c1-1500 gpurun_out/bench_reanalyze.json; tail -3 gpurun_out/bench_reanalyze.err
