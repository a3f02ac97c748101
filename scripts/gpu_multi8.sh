#!/bin/bash
# 8 GPUs of one box: the strong-scaled target at 4 and 8 ranks
cd "$(dirname "$0")/.."
TAG=${1:-r3x}
for N in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-e2e > gpurun_out/${TAG}_bench_target_n$N.json 2> gpurun_out/${TAG}_bench_target_n$N.err; echo "bench target n=$N rc=$?"
  grep '^{' gpurun_out/${TAG}_bench_target_n$N.json | head -c 700; echo; tail -2 gpurun_out/${TAG}_bench_target_n$N.err | cut -c1-300
done
