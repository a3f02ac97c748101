#!/bin/bash
# fixed cost of a stack call: the target's kernels on 1/2, 1/4 and 1/8 of its rows (what a rank of 2 / 4 / 8 GPUs gets)
cd "$(dirname "$0")/.."
TAG=${1:-s12d}
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel_ms'], d['detail']['handed_to_exact_engine_px'], d['gpu_launches'])")"
}
for h in 3194 1597 798; do
ENVV=(SGPU_PSEL_TIMING=1); run h${h}_t --height $h; tail -1 gpurun_out/${TAG}_h${h}_t.err
ENVV=(A=1); run h$h --height $h
done
