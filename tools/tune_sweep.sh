#!/bin/bash
# usage (on the GPU box): tools/tune_sweep.sh "<tune values>" "<cache log2 values>"  -- device-only bench per combination
mkdir -p gpurun_out
for cb in $2; do for t in $1; do
  HUTOKEN_B200_TUNE=$t HUTOKEN_B200_CACHE_LOG2=$cb python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline \
     > gpurun_out/sweep_t${t}_c${cb}.json 2> gpurun_out/sweep_t${t}_c${cb}.err
  echo "tune=$t cache_log2=$cb: $(python tools/bench_brief.py gpurun_out/sweep_t${t}_c${cb}.json)"
done; done
