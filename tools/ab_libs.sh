#!/bin/bash
# A/B of experimental builds of libhutoken_b200.so on the GPU box:  tools/ab_libs.sh <name> [<name> ...]
# For hutoken_b200/exp_<name>.so: the device-resident bench line with --check (8 MiB prefix vs the oracle);
# last the shipped library on the same box.  Output: gpurun_out/exp_<name>.json and one summary line each.
cd "$(dirname "$0")/.." || exit 1
line() {
    python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
    print(sys.argv[1], "MB/s", round(d["value"]), "ms/step", round(d["ms_per_step"], 3), "tile_kernel ms",
          round(d["roofline"]["kernel_ms"], 3), "frac", round(d["roofline"]["frac"], 4))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for v in "$@"; do
    HUTOKEN_B200_LIB=$PWD/hutoken_b200/exp_$v.so timeout -s KILL 300 python bench.py --no-cpu-baseline --no-e2e --no-decode --no-workloads --no-latency $AB_CHECK $AB_ARGS > gpurun_out/exp_$v.json 2> gpurun_out/exp_$v.err
    line "$v" gpurun_out/exp_$v.json
done
timeout -s KILL 300 python bench.py --no-cpu-baseline --no-e2e --no-decode --no-workloads --no-latency --check $AB_ARGS > gpurun_out/exp_base.json 2> gpurun_out/exp_base.err
line shipped gpurun_out/exp_base.json
