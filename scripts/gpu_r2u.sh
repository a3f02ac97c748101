#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r2u}
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(SGPU_PSEL_B2=1 SGPU_PSEL_STOP=4); run thread_nostore
ENVV=(SGPU_PSEL_B2=1 SGPU_PSEL_STOP=5); run thread_storeonly
ENVV=(SGPU_PSEL_B2=0 SGPU_PSEL_STOP=4); run group_nostore
ENVV=(SGPU_PSEL_B2=0 SGPU_PSEL_STOP=5); run group_storeonly
ENVV=(SGPU_PSEL_B2=1 SGPU_PSEL_STOP=3); run thread_stop3
