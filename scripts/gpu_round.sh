# one GPU round: parity tests (run under gpurun from the repo root)
set -x
TAG=${1:-r1}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/${TAG}_pytest_gpu.log
