#!/bin/bash
# usage: tools/gpu_check.sh <tag> [extra bench args]  -- GPU parity tests + a short device-only bench on one B200
tag=$1; shift
/usr/local/graft/bin/gpurun --timeout 900 -- "python -m pytest tests -m gpu -x -q 2>&1 | tail -8; python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline $* > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; python tools/bench_brief.py gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err" 2>&1 | tail -25
