#!/bin/bash
# sensitivity of the selection engine to CTAs per SM: 200 frames fit 5 CTAs; padding forces 4 and 3
cd "$(dirname "$0")/.."
TAG=${1:-r2q}
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run f200_5cta --frames 200
ENVV=(SGPU_PSEL_PAD=4096); run f200_4cta --frames 200
ENVV=(SGPU_PSEL_PAD=20000); run f200_3cta --frames 200
ENVV=(A=1); run f300_4cta
ENVV=(SGPU_PSEL_PAD=8192); run f300_3cta
