#!/bin/bash
# selection engine: parity tests, target bench (default group size and the next one), launch list + full ncu capture on a strip
cd "$(dirname "$0")/.."
TAG=${1:-r2e}
timeout 900 python -m pytest tests/test_gpu_psel.py -q > gpurun_out/${TAG}_pytest_psel.log 2>&1; echo "pytest psel rc=$?"
tail -4 gpurun_out/${TAG}_pytest_psel.log
show() { python - "$1" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1], "ms", round(d["ms_per_step"], 2), "frac", round(d["roofline"]["frac"], 4), "handed", d["detail"]["handed_to_exact_engine_px"], d["roofline"]["kernel"])
PY
}
timeout 300 python bench.py --workload target --no-e2e --steps 5 --no-cpu > gpurun_out/${TAG}_bench_target.json 2> gpurun_out/${TAG}_bench_target.err; show gpurun_out/${TAG}_bench_target.json
for lt in $ALT_LT; do SGPU_PSEL_LT=$lt timeout 300 python bench.py --workload target --no-e2e --steps 5 --no-cpu > gpurun_out/${TAG}_bench_target_lt$lt.json 2>/dev/null; show gpurun_out/${TAG}_bench_target_lt$lt.json; done
bash scripts/gpu_prof.sh target $TAG 256 stack_psel_kernel > gpurun_out/${TAG}_prof.log 2>&1; tail -1 gpurun_out/${TAG}_prof.log
