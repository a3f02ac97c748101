#!/bin/bash
# selection engine on the target: where the time goes (SGPU_PSEL_STOP), lanes per pixel, then ncu
cd "$(dirname "$0")/.."
TAG=${1:-r2k}
for s in 1 2 3 0; do
  SGPU_PSEL_STOP=$s timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_stop$s.json 2> gpurun_out/${TAG}_stop$s.err; echo "stop=$s rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_stop$s.json')); print(d['ms_per_step'], d['roofline']['frac'], d.get('detail'))")"
done
for lt in 1 3; do
  SGPU_PSEL_LT=$lt timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu > gpurun_out/${TAG}_lt$lt.json 2> gpurun_out/${TAG}_lt$lt.err; echo "lt=$lt rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_lt$lt.json')); print(d['ms_per_step'], d['roofline']['frac'], d.get('detail'))")"
done
for rej in none percentile; do
  timeout 300 python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --rej $rej > gpurun_out/${TAG}_rej$rej.json 2> gpurun_out/${TAG}_rej$rej.err; echo "rej=$rej rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/${TAG}_rej$rej.json')); print(d['ms_per_step'], d['roofline']['frac'], d.get('detail'))")"
done
timeout 900 ncu --set full --import-source on --clock-control none -k regex:stack_psel -s 3 -c 1 -o gpurun_out/${TAG}_target_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --height 1024 > gpurun_out/${TAG}_ncufull.log 2>&1; echo "ncu full rc=$?"
