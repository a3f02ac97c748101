#!/bin/bash
# N GPUs of one box: the multi-rank parity test (peer-store case included), then the strong-scaled target with the rows
# stored into rank 0's image by the kernels (peer) and with the NCCL gather after the stack (nccl)
cd "$(dirname "$0")/.."
N=${1:-2}; TAG=${2:-s12m}
timeout 900 python -m pytest tests/test_gpu_multi_rank.py -x -q > gpurun_out/${TAG}_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -5 gpurun_out/${TAG}_pytest_multi.log
for g in peer nccl; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload ${WL:-target} --gather $g --no-e2e --no-cpu > gpurun_out/${TAG}_bench_${g}_n$N.json 2> gpurun_out/${TAG}_bench_${g}_n$N.err; echo "bench $g n=$N rc=$?"
  tail -2 gpurun_out/${TAG}_bench_${g}_n$N.err | cut -c1-300
  python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_${g}_n$N.json')); print('$g', d['ms_per_step'], d['roofline']['kernel_ms'], d['value'])"
done
