#!/bin/bash
# integration glue test + the non-allocating entry point (alone, then after the rest of the suite)
cd "$(dirname "$0")/.."
TAG=${1:-r2j}
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/${TAG}_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/${TAG}_noalloc_alone.log
timeout 900 python -m pytest tests/test_gpu_integration.py -x -q > gpurun_out/${TAG}_integration.log 2>&1; echo "integration rc=$?"; tail -15 gpurun_out/${TAG}_integration.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -5 gpurun_out/${TAG}_pytest_gpu.log
