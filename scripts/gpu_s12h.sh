#!/bin/bash
# MAD clipping on the sorted engines: parity, then C2's geometry with MAD on the lane-sort engine and on the exact engine
cd "$(dirname "$0")/.."
TAG=${1:-s12h}
( time timeout 2400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_psel.py tests/test_gpu_masks.py -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run C2_mad --workload C2 --rej mad
ENVV=(SGPU_ENGINE=1); run C2_mad_exact --workload C2 --rej mad --height 1024
ENVV=(A=1); run C2_mad_h1024 --workload C2 --rej mad --height 1024
ENVV=(A=1); run target_mad --rej mad
