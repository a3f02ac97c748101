#!/bin/bash
# session 12: the whole GPU suite after the out_plane / warp-per-pixel list changes, then the target, C2 and C4 lines
cd "$(dirname "$0")/.."
TAG=${1:-s12b}
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${TAG}_pytest.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run target
ENVV=(SGPU_PSEL_TIMING=1); run target_t; tail -1 gpurun_out/${TAG}_target_t.err
ENVV=(A=1); run C2 --workload C2
ENVV=(A=1); run C4 --workload C4
