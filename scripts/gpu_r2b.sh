#!/bin/bash
# round 2, first run of the selection engine: its parity tests, then the target workload
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/r2b_pytest_psel.log 2>&1; echo "pytest psel rc=$?" 
tail -15 gpurun_out/r2b_pytest_psel.log
timeout 300 python bench.py --workload target --no-e2e --steps 3 --cpu-budget 2 > gpurun_out/r2b_bench_target.json 2> gpurun_out/r2b_bench_target.err; echo "bench rc=$?"
cat gpurun_out/r2b_bench_target.json | head -c 1500; tail -3 gpurun_out/r2b_bench_target.err
SGPU_ENGINE=5 timeout 300 python bench.py --workload target --no-e2e --steps 3 --cpu-budget 1 > gpurun_out/r2b_bench_target_lsort.json 2>/dev/null; head -c 400 gpurun_out/r2b_bench_target_lsort.json
