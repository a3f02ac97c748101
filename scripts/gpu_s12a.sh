#!/bin/bash
# session 12: group-size sweep of the selection engine on the target (SGPU_PSEL_LT), and baselines of C2 / C4 / C5-small
cd "$(dirname "$0")/.."
TAG=${1:-s12a}
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run target
ENVV=(SGPU_PSEL_LT=1); run target_lt1
ENVV=(SGPU_PSEL_LT=3); run target_lt3
ENVV=(SGPU_PSEL_LT=1 SGPU_PSEL_TIMING=1); run target_lt1_t; tail -1 gpurun_out/${TAG}_target_lt1_t.err
ENVV=(A=1); run C4 --workload C4
ENVV=(A=1); run C2 --workload C2
