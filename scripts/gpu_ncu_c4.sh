#!/bin/bash
# one full ncu capture (with source counters) of C4's lane-sort kernel on a 256-row strip
cd "$(dirname "$0")/.."
TAG=${1:-r2h}
timeout 300 ncu --set full --import-source on --clock-control none -k regex:stack_lsort -s 2 -c 1 -o gpurun_out/${TAG}_C4_lsort_full python bench.py --workload C4 --steps 1 --warmup 3 --no-e2e --no-cpu --height 256 > gpurun_out/${TAG}_ncu_c4.log 2>&1; echo "ncu rc=$?"
