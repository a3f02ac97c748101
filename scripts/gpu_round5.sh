mkdir -p gpurun_out
for alt in 0 1 2 3; do SGPU_LSORT_ALT=$alt timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --cpu-budget 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2 alt $alt ms', d['ms_per_step'], d['roofline']['frac'])"; done
for rej in none sigma; do timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --cpu-budget 1 --rej $rej 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2 rej $rej ms', d['ms_per_step'], d['roofline']['frac'])"; done
