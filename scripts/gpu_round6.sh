mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fitsread.py tests/test_gpu_parity.py -x -q -k "fits or host" > gpurun_out/r1j_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r1j_pytest.log
timeout 300 python scripts/fits_time.py > gpurun_out/r1j_fits_time.log 2>&1; echo "fits_time rc=$?"; tail -9 gpurun_out/r1j_fits_time.log
timeout 600 python bench.py --steps 3 --warmup 3 --cpu-budget 1 2>/dev/null | tail -c 500
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1j_pytest_gpu.log 2>&1; echo "pytest all rc=$?"; tail -3 gpurun_out/r1j_pytest_gpu.log
