# profile one workload: launch list of the stack kernels + one full capture of the top kernel
set -x
WL=${1:-C2}; TAG=${2:-r2}; H=${3:-512}; KERN=${4:-stack_reg_kernel}
mkdir -p gpurun_out
timeout 600 python bench.py --workload $WL --steps 3 --warmup 3 --no-e2e --no-cpu --height $H $EXTRA > gpurun_out/${TAG}_${WL}_h${H}.json 2>&1; echo "rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:stack_ -c 24 --csv --log-file gpurun_out/${TAG}_launches_${WL}.csv python bench.py --workload $WL --steps 2 --warmup 3 --no-e2e --height $H --no-cpu $EXTRA > gpurun_out/${TAG}_ncu_${WL}.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$KERN -s 3 -c 1 -o gpurun_out/${TAG}_${WL}_full python bench.py --workload $WL --steps 2 --warmup 3 --no-e2e --height $H --no-cpu $EXTRA > gpurun_out/${TAG}_ncufull_${WL}.log 2>&1; echo "ncu full rc=$?"
