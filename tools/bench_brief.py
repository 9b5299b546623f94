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
    for name, w in [("main", d)] + list((d.get("workloads") or {}).items()):
        dec = w.get("decode") if isinstance(w, dict) else None
        if dec and dec.get("value"):
            one = dec.get("single_document", {})
            print(f"decode[{name}] {dec['value']:.0f} MB/s {dec['ms_per_step']:.3f} ms path {dec.get('path')} round_trip "
                  f"{dec.get('round_trip_equals_input')} oracle {dec.get('first_docs_equal_oracle')}; one document "
                  f"{one.get('value', 0):.0f} MB/s x{one.get('ms_vs_batch', 0):.2f} {one.get('path')}")
        elif dec:
            print(f"decode[{name}] {dec}")
    for name, w in (d.get("workloads") or {}).items():
        print(f"workload {name}: {w.get('value', 0):.0f} MB/s parity {w.get('parity', {}).get('equal_to_oracle')}")
