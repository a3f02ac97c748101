#!/bin/bash
# median stacks on the selection engine with float planning sums: parity, then C3 and the target (unchanged path)
cd "$(dirname "$0")/.."
TAG=${1:-s12p}
( time timeout 2400 python -m pytest tests/test_gpu_psel.py tests/test_gpu_parity.py -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -1
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run C3 --workload C3
ENVV=(A=1); run target
