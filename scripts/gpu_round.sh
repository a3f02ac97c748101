# one GPU round: parity tests, bench lines (run under gpurun from the repo root)
set -x
TAG=${1:-r1}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/${TAG}_pytest_gpu.log
for mb in 3 2; do
SGPU_TSORT_MINB=$mb timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --cpu-budget 2 > gpurun_out/${TAG}_bench_C2_mb$mb.json 2> gpurun_out/${TAG}_bench_C2_mb$mb.err; echo "bench rc=$?"
cat gpurun_out/${TAG}_bench_C2_mb$mb.json; tail -n 5 gpurun_out/${TAG}_bench_C2_mb$mb.err
done
