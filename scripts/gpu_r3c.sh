#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r3c}
timeout 1200 python -m pytest tests/test_gpu_psel.py tests/test_gpu_parity.py tests/test_gpu_fitsread.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
SGPU_PSEL_CHUNK=1000 timeout 1200 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/${TAG}_pytest_chunk.log 2>&1; echo "pytest chunked rc=$?"; tail -2 gpurun_out/${TAG}_pytest_chunk.log
bash scripts/gpu_launchlist.sh ${TAG}_C3 --workload C3
bash scripts/gpu_launchlist.sh ${TAG}_f100 --frames 100
bash scripts/gpu_launchlist.sh ${TAG}_target
