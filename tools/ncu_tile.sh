#!/bin/bash
# ncu --set full capture of one tile_kernel launch of the device-resident bench leg:  tools/ncu_tile.sh <tag> [lib.so]
cd "$(dirname "$0")/.." || exit 1
tag=$1
[ -n "$2" ] && export HUTOKEN_B200_LIB=$PWD/$2
ARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-decode --no-workloads --no-latency"
timeout -s KILL 300 python bench.py $ARGS > gpurun_out/ncu_${tag}_plain.json 2> gpurun_out/ncu_${tag}_plain.err || { echo "plain run failed"; exit 1; }
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 1 -c 1 -f -o gpurun_out/ncu_$tag \
    python bench.py $ARGS > gpurun_out/ncu_${tag}.log 2>&1
ls -la gpurun_out/ncu_$tag.ncu-rep
