#!/bin/bash
cd "$(dirname "$0")/.."
for e in none 1; do
SGPU_NO_PEER=$e timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-e2e --no-cpu --height 2048 > gpurun_out/s12g_$e.json 2> gpurun_out/s12g_$e.err; echo "rc=$?"
grep "falling back" gpurun_out/s12g_$e.err
python -c "import json; d=json.load(open('gpurun_out/s12g_$e.json')); print(d['ms_per_step'], d['config']['sharding'][:120])"
done
