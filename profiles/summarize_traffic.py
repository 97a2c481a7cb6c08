#!/usr/bin/env python
"""Per-kernel sums of gpu__time_duration / dram bytes from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum --csv` log.   python profiles/summarize_traffic.py gpurun_out/x.csv [skip_first_n_launches_of_ce_bwd]"""
import collections, csv, re, sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
per = collections.OrderedDict()
ids = {}
for r in csv.DictReader(lines):
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("bsa::", "").replace("(anonymous namespace)::", "")[:72]
    rec = per.setdefault(name, {"n": set(), "t": 0.0, "rd": 0.0, "wr": 0.0})
    rec["n"].add(r["ID"])
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}.get(unit, 1)
    if r["Metric Name"] == "gpu__time_duration.sum": rec["t"] += v * scale
    elif r["Metric Name"] == "dram__bytes_read.sum": rec["rd"] += v * scale
    elif r["Metric Name"] == "dram__bytes_write.sum": rec["wr"] += v * scale
for name, rec in per.items():
    print(f"{name:72s} x{len(rec['n']):4d}  time {rec['t']:9.3f} ms  dram read {rec['rd'] / 1e6:10.1f} MB  write {rec['wr'] / 1e6:10.1f} MB")
