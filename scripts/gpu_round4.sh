set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "host" > gpurun_out/r1i_pytest_host.log 2>&1; echo "pytest host rc=$?"; tail -4 gpurun_out/r1i_pytest_host.log
timeout 900 python bench.py > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err; echo "bench rc=$?"; tail -c 700 gpurun_out/r1i_bench.json
for nb in 1 4 16; do SGPU_HOST_BLOCKS=$nb timeout 900 python bench.py --steps 3 --warmup 3 --cpu-budget 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('blocks $nb e2e ms', d['e2e']['ms_per_step'])"; done
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r1i_pytest_gpu.log
