#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r3f}
SGPU_PSEL_TIMING=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_t.json 2> gpurun_out/${TAG}_t.err; tail -4 gpurun_out/${TAG}_t.err; python -c "import json; d=json.load(open('gpurun_out/${TAG}_t.json')); print(d['ms_per_step'])"
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_a.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/${TAG}_a.json')); print('plain', d['ms_per_step'])"
SGPU_PSEL_TIMING=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu --workload C2 > gpurun_out/${TAG}_c2.json 2> gpurun_out/${TAG}_c2.err; tail -2 gpurun_out/${TAG}_c2.err; python -c "import json; d=json.load(open('gpurun_out/${TAG}_c2.json')); print(d['ms_per_step'])"
