"""prints the few numbers of a bench.py JSON line one looks at while optimising"""
import json, sys
for line in open(sys.argv[1]):
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    r = d.get("roofline", {})
    print(f"value {d['value']:.0f} {d['unit']}  ms/step {d['ms_per_step']:.3f}  kernel_ms {r.get('kernel_ms', 0):.3f}  "
          f"frac {r.get('frac', 0):.4f}  stats {d.get('stats_last_step')}  e2e {d.get('e2e', {}).get('value')}")
