#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r2r}
timeout 1200 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest.log
SGPU_PSEL_B2=1 timeout 1200 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/${TAG}_pytest_b2.log 2>&1; echo "pytest b2=1 rc=$?"; tail -2 gpurun_out/${TAG}_pytest_b2.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(SGPU_PSEL_B2=0); run b2group
ENVV=(SGPU_PSEL_B2=1); run b2thread
ENVV=(SGPU_PSEL_B2=1); run b2thread_none --rej none
ENVV=(SGPU_PSEL_B2=1); run b2thread_C3 --workload C3
