#!/bin/bash
# SIGMEDIAN on the sorted engines: the whole GPU suite, then the target and C2 / C4 (no regression), C2's geometry with sigmedian
cd "$(dirname "$0")/.."
TAG=${1:-s12i}
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -2
run() { local name=$1; shift
  env "${ENVV[@]}" timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_$name.json 2> gpurun_out/${TAG}_$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_$name.json')); print(d['ms_per_step'], round(d['roofline']['frac'],4), d['roofline']['kernel'], d['detail']['handed_to_exact_engine_px'])")"
}
ENVV=(A=1); run target
ENVV=(A=1); run C2 --workload C2
ENVV=(A=1); run C4 --workload C4
ENVV=(A=1); run C2_sigmedian --workload C2 --rej sigmedian
ENVV=(SGPU_ENGINE=1); run C2_sigmedian_exact --workload C2 --rej sigmedian --height 1024
