#!/bin/bash
# N GPUs of one box: the multi-rank parity test, then the strong-scaled bench (target; C2 with sharded statistics)
cd "$(dirname "$0")/.."
N=${1:-2}; TAG=${2:-r2m}
timeout 900 python -m pytest tests/test_gpu_multi_rank.py -x -q > gpurun_out/${TAG}_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -5 gpurun_out/${TAG}_pytest_multi.log
for wl in $WLS; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $wl $EXTRA > gpurun_out/${TAG}_bench_${wl}_n$N.json 2> gpurun_out/${TAG}_bench_${wl}_n$N.err; echo "bench $wl n=$N rc=$?"
  tail -2 gpurun_out/${TAG}_bench_${wl}_n$N.err | cut -c1-300; head -c 1800 gpurun_out/${TAG}_bench_${wl}_n$N.json; echo
done
