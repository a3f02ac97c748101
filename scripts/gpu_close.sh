#!/bin/bash
# closing check of the round: the whole GPU suite, smoke(), the default bench line
cd "$(dirname "$0")/.."
TAG=${1:-r2z}
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest_gpu.log | tail -1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
timeout 1200 python bench.py > gpurun_out/${TAG}_bench_target.json 2> gpurun_out/${TAG}_bench_target.err; echo "bench rc=$?"
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_target.json')); print(d['ms_per_step'], d['roofline']['frac'], d['roofline']['traffic'], d['e2e']['value'], d['cpu_baseline']['value'], d['gpu_launches'], d['clocks'])"
