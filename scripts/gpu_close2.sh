#!/bin/bash
# closing check after the GESDT change: whole GPU suite, smoke(), default bench line, C4 / C2 lines, one ncu capture of C4's kernel
cd "$(dirname "$0")/.."
TAG=${1:-r2y}
bash scripts/gpu_close.sh $TAG
for w in C4 C2; do
timeout 200 python bench.py --workload $w --no-e2e --no-cpu > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err; echo "bench $w rc=$?"
python -c "import json; d=json.load(open('gpurun_out/${TAG}_bench_$w.json')); print('$w', d['ms_per_step'], d['roofline']['frac'], d['gpu_launches'], d['detail'])"
done
timeout 120 ncu --set full --import-source on --clock-control none -k regex:stack_lsort -s 2 -c 1 -o gpurun_out/${TAG}_C4_lsort_full python bench.py --workload C4 --steps 1 --warmup 3 --no-e2e --no-cpu --height 256 > gpurun_out/${TAG}_ncu_c4.log 2>&1; echo "ncu rc=$?"
