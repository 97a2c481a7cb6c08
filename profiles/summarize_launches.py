#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.

    python profiles/summarize_launches.py gpurun_out/launches.csv [first_id last_id]
"""
import csv, re, sys, collections

def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("bsa::<unnamed>::", "").replace("bsa::umma::", "").replace("bsa::", "")
    name = re.sub(r"at::native::.*?(\w+_kernel\w*|\w+Functor\w*).*", r"at::\1", name)
    return name[:110]

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((int(r["ID"]), short(r["Kernel Name"]), float(r["Metric Value"].replace(",", "")), r["Grid Size"], r["Block Size"]))
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 10**9
rows = [r for r in rows if lo <= r[0] <= hi]
agg = collections.OrderedDict()
for _, n, ns, g, b in rows:
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1; a[1] += ns
tot = sum(v[1] for v in agg.values())
print(f"launches {len(rows)}  total {tot/1e6:.3f} ms (cold-cache, serialised: compare shares)")
for n, (c, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ns/1e6:9.3f} ms {100*ns/tot:5.1f}%  x{c:<4d} {n}")
