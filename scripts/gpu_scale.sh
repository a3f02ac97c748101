#!/bin/bash
# strong-scaled target on N GPUs (rows stored into rank 0's image by the kernels), bench line kept
cd "$(dirname "$0")/.."
N=${1:-2}; TAG=${2:-r2f}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-e2e --no-cpu > gpurun_out/${TAG}_bench_target_n$N.json 2> gpurun_out/${TAG}_bench_target_n$N.err; echo "bench n=$N rc=$?"
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_target_n$N.json')); print(d['n_gpus'], d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['detail'])"
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi_rank.py -x -q 2>&1 | tail -2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload C2 --no-e2e --no-cpu > gpurun_out/${TAG}_bench_C2_n$N.json 2> gpurun_out/${TAG}_bench_C2_n$N.err; echo "bench C2 n=$N rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_C2_n$N.json')); print(d['n_gpus'], d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['detail'])"
fi
