# round-end check on the GPU box: new FITS reader parity, smoke, default bench, then the full parity suite
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fitsread.py -x -q > gpurun_out/r1f_pytest_fits.log 2>&1; echo "fits rc=$?"
tail -5 gpurun_out/r1f_pytest_fits.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r1f_smoke.log
timeout 900 python bench.py > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r1f_bench.json
timeout 300 python scripts/fits_time.py > gpurun_out/r1f_fits_time.log 2>&1; echo "fits_time rc=$?"; tail -8 gpurun_out/r1f_fits_time.log
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_fitsread.py > gpurun_out/r1f_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r1f_pytest_gpu.log
