#!/bin/bash
# GESDT cases of the GPU suite + the C4 bench line (lane-sort engine, Grubbs loop)
cd "$(dirname "$0")/.."
TAG=${1:-r2g}
( time timeout 240 python -m pytest tests/test_gpu_parity.py tests/test_gpu_masks.py tests/test_gpu_psel.py -m gpu -x -q -k "gesdt or GESDT or engines or fuzz" ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|error" gpurun_out/${TAG}_pytest.log | tail -3
timeout 200 python bench.py --workload C4 --no-e2e --no-cpu > gpurun_out/${TAG}_bench_C4.json 2> gpurun_out/${TAG}_bench_C4.err; echo "bench rc=$?"
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_C4.json')); print(d['ms_per_step'], d['roofline']['frac'], d['gpu_launches'], d.get('detail'))"
