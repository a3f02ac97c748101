#!/bin/bash
# where C4's time goes: the same geometry through the lane-sort engine without the Grubbs loop
cd "$(dirname "$0")/.."
for r in none sigma; do
SGPU_ENGINE=5 timeout 200 python bench.py --workload C4 --rej $r --no-e2e --no-cpu > gpurun_out/c4split_$r.json 2> gpurun_out/c4split_$r.err
python -c "import json; d=json.load(open('gpurun_out/c4split_$r.json')); print('$r', d['ms_per_step'], d['gpu_launches'], d['detail']['handed_to_exact_engine_px'])"
done
