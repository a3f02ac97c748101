#!/bin/bash
cd "$(dirname "$0")/.."
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_multi_rank.py -x -q 2>&1 | tail -4
for st in frames reduce; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload C2 --stats $st --no-e2e --no-cpu > gpurun_out/s12k_C2_${st}_n$N.json 2> gpurun_out/s12k_C2_${st}_n$N.err; echo "bench C2 $st n=$N rc=$?"; grep -i "error\|Traceback" -A5 gpurun_out/s12k_C2_${st}_n$N.err | head -20
python -c "import json; d=json.load(open('gpurun_out/s12k_C2_${st}_n$N.json')); print(d['n_gpus'], d['ms_per_step'], d['value'], d['detail'])"
done
