#!/bin/bash
# GPU box: parity tests that exercise the deferred-pretoken kernels, then the heavy-tail workload alone (both shapes)
# and the launch list of its kernels.  usage: tools/heavy_check.sh
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_merges.py tests/test_config0.py -m gpu -x -q 2>&1 | tail -4
python tools/heavy_probe.py > gpurun_out/hp.log 2>&1; tail -3 gpurun_out/hp.log | cut -c1-60
python tools/heavy_probe.py llama > gpurun_out/hpl.log 2>&1; tail -2 gpurun_out/hpl.log | cut -c1-60
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_hp.csv python tools/heavy_probe.py > gpurun_out/ncu_hp.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches_hp.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[-9:-3]: print(r[4][:40], r[7], r[8], r[-1])
PY
