#!/bin/bash
# the multi-GPU configs BASELINE.json names, strong-scaled: given workloads on N GPUs
cd "$(dirname "$0")/.."
N=${1:-8}; shift
for wl in "$@"; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload $wl --no-e2e --no-cpu > gpurun_out/r2f_bench_${wl}_n$N.json 2> gpurun_out/r2f_bench_${wl}_n$N.err; echo "bench $wl n=$N rc=$?"; grep -i "Traceback" -A8 gpurun_out/r2f_bench_${wl}_n$N.err | head -12
  python -c "import json; d=json.load(open('gpurun_out/r2f_bench_${wl}_n$N.json')); print(d['n_gpus'], d['ms_per_step'], d['roofline']['kernel_ms'], d['value'], d['detail'])"
done
