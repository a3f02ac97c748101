#!/bin/bash
# one full ncu capture of the selection-engine kernel on a 1024-row strip of the target
cd "$(dirname "$0")/.."
TAG=${1:-r2p}
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel -s 3 -c 1 -o gpurun_out/${TAG}_target_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull.log 2>&1; echo "ncu full rc=$?"
