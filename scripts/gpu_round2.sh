set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fitsread.py tests/test_gpu_simple.py -x -q > gpurun_out/r1g_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r1g_pytest.log
timeout 300 python scripts/fits_time.py > gpurun_out/r1g_fits_time.log 2>&1; echo "fits_time rc=$?"; tail -8 gpurun_out/r1g_fits_time.log
(timeout 300 python scripts/simple_time.py 100 4176 6248 f32; timeout 300 python scripts/simple_time.py 100 4176 6248 u16; SGPU_SIMPLE_SCALAR=1 timeout 300 python scripts/simple_time.py 100 4176 6248 u16) > gpurun_out/r1g_simple_time.log 2>&1; tail -20 gpurun_out/r1g_simple_time.log
