#!/bin/bash
# A/B of experimental builds of libhutoken_b200.so on the GPU box:  tools/ab_libs.sh <name> [<name> ...]
# For hutoken_b200/exp_<name>.so: the C-ABI parity tests, then the device-resident bench line with --check;
# last the shipped library on the same box.  Output: gpurun_out/exp_<name>.json and one summary line each.
cd "$(dirname "$0")/.." || exit 1
line() {
    python - "$1" "$2" <<'PY'
import json, sys
d = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
print(sys.argv[1], "MB/s", round(d["value"]), "ms/step", round(d["ms_per_step"], 3), "tile_kernel ms",
      round(d["roofline"]["kernel_ms"], 3), "frac", round(d["roofline"]["frac"], 4), "check", d.get("check"))
PY
}
for v in "$@"; do
    export HUTOKEN_B200_LIB=$PWD/hutoken_b200/exp_$v.so
    python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
    python bench.py --no-cpu-baseline --no-e2e --no-decode --check > gpurun_out/exp_$v.json 2> gpurun_out/exp_$v.err
    line "$v" gpurun_out/exp_$v.json
done
unset HUTOKEN_B200_LIB
python bench.py --no-cpu-baseline --no-e2e --no-decode > gpurun_out/exp_base.json 2> gpurun_out/exp_base.err
line shipped gpurun_out/exp_base.json
