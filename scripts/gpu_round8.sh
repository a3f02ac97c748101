mkdir -p gpurun_out
for wl in target C3 C4 C1; do
timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --cpu-budget 1 2>/dev/null | tail -1 > gpurun_out/r1l_bench_$wl.json
python -c "import sys,json; d=json.loads(open('gpurun_out/r1l_bench_$wl.json').read()); print('$wl', round(d['ms_per_step'],3), 'ms', round(d['roofline']['achieved']), 'GB/s frac', round(d['roofline']['frac'],4), 'value', d['value'])"
done
