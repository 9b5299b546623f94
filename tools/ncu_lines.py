"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by source file:line.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python tools/ncu_lines.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
hdr = None
agg = []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or r[0] in ("Function Name",) or not r[0].isdigit():
        continue
    d = dict(zip(hdr[4:], r[4:]))
    def num(k):
        try:
            return float(d.get(k, "0") or 0)
        except ValueError:
            return 0.0
    agg.append((cur_file, int(r[0]), r[1].strip()[:70], num("# Samples"), num("Instructions Executed"),
                num("stall_barrier"), num("stall_long_sb"), num("stall_wait"), num("stall_short_sb")))
tot_s = sum(a[3] for a in agg) or 1
tot_i = sum(a[4] for a in agg) or 1
print(f"total samples {tot_s:.0f}, total warp instructions {tot_i:.3e}")
print("--- by samples")
for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print(f"{a[0]}:{a[1]:<5} smp {100*a[3]/tot_s:5.1f}%  inst {100*a[4]/tot_i:5.1f}%  bar {a[5]:7.0f} lsb {a[6]:7.0f} wait {a[7]:6.0f} ssb {a[8]:6.0f} | {a[2]}")
print("--- by instructions")
for a in sorted(agg, key=lambda a: -a[4])[:top]:
    print(f"{a[0]}:{a[1]:<5} inst {100*a[4]/tot_i:5.1f}%  smp {100*a[3]/tot_s:5.1f}% | {a[2]}")
