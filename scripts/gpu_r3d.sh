#!/bin/bash
cd "$(dirname "$0")/.."
TAG=${1:-r3d}
timeout 1200 python -m pytest tests/test_gpu_psel.py tests/test_gpu_parity.py -x -q -k "psel or auto" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
SGPU_PSEL_CHUNK=1000 timeout 1200 python -m pytest tests/test_gpu_psel.py -x -q > gpurun_out/${TAG}_pytest_chunk.log 2>&1; echo "pytest chunked rc=$?"; tail -2 gpurun_out/${TAG}_pytest_chunk.log
bash scripts/gpu_launchlist.sh ${TAG}_target
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel_kernel -s 3 -c 1 -o gpurun_out/${TAG}_k1_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull1.log 2>&1; echo "ncu k1 rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel_finish -s 3 -c 1 -o gpurun_out/${TAG}_k2_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull2.log 2>&1; echo "ncu k2 rc=$?"
