#!/bin/bash
# usage: tools/launch_list.sh <tag> <skip> <count> [bench args] -- per-launch device times of one bench step under ncu
tag=$1; skip=$2; cnt=$3; shift 3
/usr/local/graft/bin/gpurun --timeout 900 -- "python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline $* > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s $skip -c $cnt --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline $* > gpurun_out/ncu.log 2>&1; tail -2 gpurun_out/ncu.log" 2>&1 | tail -2
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_$tag.csv')) if len(r)>10 and r[0].isdigit()]
tot={}
for r in rows:
    print(r[4][:40], r[7], r[8], r[-1])
    k=r[4].split('(')[0]; tot[k]=tot.get(k,0)+float(r[-1])
print({k:round(v/1e6,3) for k,v in tot.items()})
PY
