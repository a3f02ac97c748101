#!/bin/bash
# launch list (kernel durations) of one bench command: TAG then bench arguments
cd "$(dirname "$0")/.."
TAG=$1; shift
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:stack_ -c 60 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/${TAG}_launches.csv')) if len(r)>5]
h=rows[0]; ik=h.index('Kernel Name'); iv=h.index('Metric Value')
for r in rows[1:][-9:]: print(r[ik][:60], r[iv])
PY
