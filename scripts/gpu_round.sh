set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu.log 2>&1; echo "pytest rc=$?" 
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r1_bench_C2.json 2> gpurun_out/r1_bench_C2.err; echo "bench rc=$?"
timeout 600 python bench.py --workload target --steps 3 --warmup 3 --no-e2e --cpu-budget 5 > gpurun_out/r1_bench_target.json 2> gpurun_out/r1_bench_target.err; echo "bench target rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches_C2.csv python bench.py --steps 2 --warmup 3 --no-e2e --height 512 --cpu-budget 1 > gpurun_out/r1_ncu_C2.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_bench_C2.json gpurun_out/r1_bench_target.json
