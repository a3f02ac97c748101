#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-s12e}
( time timeout 2400 python -m pytest tests/test_gpu_psel.py -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel_ms'], d['detail']['handed_to_exact_engine_px'], d['gpu_launches'])")"
}
ENVV=(A=1); run target
for h in 3194 798; do
ENVV=(A=1); run h$h --height $h
done
