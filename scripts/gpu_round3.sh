set -x
mkdir -p gpurun_out
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1h_bench_ref.json 2> gpurun_out/r1h_bench_ref.err; echo "ref rc=$?"; tail -c 900 gpurun_out/r1h_bench_ref.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r1h_bench_n2.json 2> gpurun_out/r1h_bench_n2.err; echo "n2 rc=$?"; tail -c 1800 gpurun_out/r1h_bench_n2.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'simple_stack|fits_decode' -c 40 --csv --log-file gpurun_out/r1h_launches_simple_fits.csv python scripts/simple_fits_once.py > gpurun_out/r1h_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r1h_ncu.log
