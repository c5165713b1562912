"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
python tools/ncu_summary.py launches.csv "<command that was profiled>" > summary.json"""
import csv
import io
import json
import re
import sys

text = open(sys.argv[1], errors="replace").read()
start = text.find('"ID"')
rows = list(csv.DictReader(io.StringIO(text[start:])))
by = {}
for r in rows:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("void ", "").replace("unnamed>::", "")
    name = name.replace("rdl::<unnamed>::", "").replace("rdl::(anonymous namespace)::", "").replace("rdl::", "")
    ns = float(r["Metric Value"].replace(",", ""))
    if r.get("Metric Unit", "ns") in ("us", "usecond"):
        ns *= 1e3
    elif r.get("Metric Unit") in ("ms", "msecond"):
        ns *= 1e6
    k = by.setdefault(name, [0.0, 0])
    k[0] += ns
    k[1] += 1
total = sum(v[0] for v in by.values())
out = {"command": sys.argv[2] if len(sys.argv) > 2 else "", "total_ms": total / 1e6,
       "launches": sum(v[1] for v in by.values()),
       "by_kernel": [{"kernel": k, "total_ms": v[0] / 1e6, "share": v[0] / total, "launches": v[1],
                      "avg_us": v[0] / v[1] / 1e3}
                     for k, v in sorted(by.items(), key=lambda kv: -kv[1][0])]}
print(json.dumps(out, indent=1))
