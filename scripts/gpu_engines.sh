# compare the engines on one workload (SGPU_ENGINE override): bench lines only
WL=${1:-C2}; TAG=${2:-r5}; shift 2
mkdir -p gpurun_out
for e in "$@"; do
SGPU_ENGINE=$e timeout 600 python bench.py --workload $WL --steps 3 --warmup 3 --no-e2e --cpu-budget 1 > gpurun_out/${TAG}_${WL}_e$e.json 2> gpurun_out/${TAG}_${WL}_e$e.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_${WL}_e$e.json").read().strip().splitlines()[-1])
print("$WL engine $e: ms", round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],4), "handed", d["config"]["handed_to_exact_engine_px"])
PY
done
