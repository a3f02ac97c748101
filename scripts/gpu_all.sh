# full round: all GPU tests, then one bench line per workload (device-resident; e2e only for the default)
TAG=${1:-r9}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "default rc=$?"; cat gpurun_out/${TAG}_bench_default.json
for wl in C1 C3 C4 C5 target; do
timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --cpu-budget 3 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "$wl rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench_$wl.json").read().strip().splitlines()[-1])
    print("$wl: ms", round(d["ms_per_step"],3), "value", "%.4g"%d["value"], "frac", round(d["roofline"]["frac"],4), "kernel", d["roofline"]["kernel"], "handed", d["config"]["handed_to_exact_engine_px"], "cpu", "%.3g"%d["cpu_baseline"]["value"])
except Exception as e: print("$wl: failed", e)
PY
done
