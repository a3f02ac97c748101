#!/bin/bash
# round 2, closing measurement set (1 GPU): full GPU suite, smoke, default bench (target, e2e + cpu_baseline), reference arm,
# device-resident lines of C1-C5, launch list of the target, full ncu captures (target's two kernels, C4's lane-sort kernel)
cd "$(dirname "$0")/.."
TAG=${1:-r2f}
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 1200 python bench.py > gpurun_out/${TAG}_bench_target.json 2> gpurun_out/${TAG}_bench_target.err; echo "bench rc=$?"; tail -2 gpurun_out/${TAG}_bench_target.err; head -c 300 gpurun_out/${TAG}_bench_target.json; echo
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_target_reference_arm.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"; head -c 300 gpurun_out/${TAG}_bench_target_reference_arm.json; echo
for wl in C1 C2 C3 C4 C5; do
  timeout 900 python bench.py --workload $wl --no-e2e > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "$wl rc=$? $(python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_$wl.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail'], d['cpu_baseline']['value'])")"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:stack_ -c 24 --csv --log-file gpurun_out/${TAG}_launches_target.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_ncu_ll.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel_kernel -s 3 -c 1 -o gpurun_out/${TAG}_target_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel_finish -s 3 -c 1 -o gpurun_out/${TAG}_target_finish_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull2.log 2>&1; echo "ncu full (finish kernel) rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_lsort_kernel -s 3 -c 1 -o gpurun_out/${TAG}_C4_lsort_full python bench.py --workload C4 --steps 2 --warmup 3 --no-e2e --no-cpu --height 256 > gpurun_out/${TAG}_ncufull3.log 2>&1; echo "ncu full (C4 lane-sort) rc=$?"
ls -la gpurun_out/*.ncu-rep
