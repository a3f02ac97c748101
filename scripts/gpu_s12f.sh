#!/bin/bash
# GESDT loop with table reciprocals and the test in squares: parity, then C4
cd "$(dirname "$0")/.."
TAG=${1:-s12f}
( time timeout 2400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_psel.py tests/test_gpu_integration.py -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run C4 --workload C4
ENVV=(A=1); run C4_none --workload C4 --rej none
