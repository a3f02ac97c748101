#!/bin/bash
# selection engine: parity tests, target bench, launch list + full ncu capture on a strip
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/r2d_pytest_psel.log 2>&1; echo "pytest psel rc=$?"
tail -5 gpurun_out/r2d_pytest_psel.log
timeout 300 python bench.py --workload target --no-e2e --steps 5 --cpu-budget 1 > gpurun_out/r2d_bench_target.json 2> gpurun_out/r2d_bench_target.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2d_bench_target.json"))
print("target: ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "handed", d["config"]["handed_to_exact_engine_px"], d["roofline"]["kernel"])
PY
bash scripts/gpu_prof.sh target r2d 256 stack_psel_kernel > gpurun_out/r2d_prof.log 2>&1; tail -2 gpurun_out/r2d_prof.log
