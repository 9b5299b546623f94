#!/bin/bash
# tile_kernel time against the size of the pretoken cache, on a corpus whose distinct pretokens leave even the smallest
# table lightly loaded:  tools/cache_sweep.sh <MiB of corpus> <log2 slots> ...
cd "$(dirname "$0")/.." || exit 1
mb=$1; shift
for b in "$@"; do
    HUTOKEN_B200_CACHE_LOG2=$b timeout -s KILL 300 python bench.py --mb $mb --no-cpu-baseline --no-e2e --no-decode --no-workloads --no-latency --check > gpurun_out/sweep_${mb}_$b.json 2> gpurun_out/sweep_${mb}_$b.err
    python - $mb $b gpurun_out/sweep_${mb}_$b.json <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[3]) if l.startswith("{")][-1])
    s = d.get("stats_last_step", {})
    print("MiB", sys.argv[1], "log2 slots", sys.argv[2], "tile_kernel ms", round(d["roofline"]["kernel_ms"], 3), "ms per GiB",
          round(d["roofline"]["kernel_ms"] * 1024 / int(sys.argv[1]), 3), "misses", s.get("lane_merges"), "check", d.get("check"))
except Exception as e:
    print(sys.argv[1], sys.argv[2], "FAILED", e)
PY
done
