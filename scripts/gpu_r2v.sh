#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r2v}
timeout 1200 python -m pytest tests/test_gpu_psel.py tests/test_gpu_parity.py -x -q -k "psel or auto" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run target
ENVV=(SGPU_PSEL_STOP=1); run stop1
ENVV=(SGPU_PSEL_STOP=2); run stop2
ENVV=(SGPU_PSEL_STOP=3); run stop3
ENVV=(A=1); run none --rej none
ENVV=(A=1); run C3 --workload C3
ENVV=(A=1); run C1 --workload C1
