A="--steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-latency --no-decode --llama-mb 64"
for v in old new old new; do
  if [ $v = old ]; then export HUTOKEN_B200_LIB=$PWD/hutoken_b200/exp_old.so; else unset HUTOKEN_B200_LIB; fi
  python bench.py $A > gpurun_out/ht_$v.json 2> gpurun_out/ht_$v.err
  python - $v <<'PY'
import json,sys
d=json.loads([l for l in open(f'gpurun_out/ht_{sys.argv[1]}.json') if l.startswith('{')][-1])
w=d['workloads']
print(sys.argv[1], 'heavy', round(w['heavy_tail_mix']['value']), w['heavy_tail_mix']['ms_per_step'], w['heavy_tail_mix']['parity']['equal_to_oracle'], 'cfg4 ms', w['configs[4] long_runs_no_limit']['ms_per_step'], w['configs[4] long_runs_no_limit']['parity']['equal_to_oracle'])
PY
done
