#!/bin/bash
# A/B of experimental builds on the decode leg:  tools/ab_decode.sh <name> [<name> ...]   (run on the GPU box)
# For hutoken_b200/exp_<name>.so and then the shipped library: `decode` object of the bench line (round trip checked).
cd "$(dirname "$0")/.." || exit 1
A="--steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-workloads --no-latency $AB_ARGS"
line() {
    python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])["decode"]
    print(sys.argv[1], "decode MB/s", round(d["value"]), "ms", round(d["ms_per_step"], 3), "same", d["round_trip_equals_input"], d.get("path"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for v in "$@"; do
    HUTOKEN_B200_LIB=$PWD/hutoken_b200/exp_$v.so timeout -s KILL 300 python bench.py $A > gpurun_out/dexp_$v.json 2> gpurun_out/dexp_$v.err
    line "$v" gpurun_out/dexp_$v.json
done
timeout -s KILL 300 python bench.py $A > gpurun_out/dexp_base.json 2> gpurun_out/dexp_base.err
line shipped gpurun_out/dexp_base.json
