"""Sum an ncu launch-list CSV (gpu__time_duration.sum) by kernel name: python tools/sum_launches.py file.csv [filter]"""
import csv, sys, collections, re
rows = [r for r in csv.reader(l for l in open(sys.argv[1], errors="replace") if l.startswith('"'))]
hdr = rows[0]; ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
t = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if len(sys.argv) > 2 and sys.argv[2] not in r[ki]: continue
    v = float(r[vi].replace(",", "")); v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(r[ui], 1.0)
    name = re.sub(r"\(.*", "", r[ki]); t[name][0] += 1; t[name][1] += v
tot = sum(v[1] for v in t.values())
for k, v in sorted(t.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:40s} {v[0]:5d} launches {v[1]:10.1f} us {100 * v[1] / tot:5.1f} %")
print(f"{'total':40s} {sum(v[0] for v in t.values()):5d} launches {tot:10.1f} us")
