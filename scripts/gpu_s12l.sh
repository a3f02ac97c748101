#!/bin/bash
# statistics of C2's 100 planes against the number of worker contexts
cd "$(dirname "$0")/.."
for wk in 1 2 3 4 6 8; do
SGPU_STATS_WORKERS=$wk timeout 300 python bench.py --workload C2 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/s12l_$wk.json 2> gpurun_out/s12l_$wk.err
python -c "import json; d=json.load(open('gpurun_out/s12l_$wk.json')); print($wk, d['detail']['stats_ms'])"
done
