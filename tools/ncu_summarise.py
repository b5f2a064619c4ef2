"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name the summed device time,
share and launch count over launches [skip, skip+count).  usage: ncu_summarise.py launches.csv [skip] [count]"""
import csv
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
count = int(sys.argv[3]) if len(sys.argv) > 3 else 10 ** 9
rows = []
with open(path, newline="") as f:
    lines = [ln for ln in f if not ln.startswith("==")]
hdr = next(i for i, ln in enumerate(lines) if ln.startswith('"ID"'))
for r in csv.DictReader(lines[hdr:]):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ns = v * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
    name = re.sub(r"^(void )?(vb::)?(\(anonymous namespace\)::)?", "", r["Kernel Name"])
    name = re.sub(r"\(.*$", "", name)
    rows.append((name, ns))
rows = rows[skip:skip + count]
agg = OrderedDict()
for n, ns in rows:
    a = agg.setdefault(n, [0.0, 0])
    a[0] += ns
    a[1] += 1
tot = sum(a[0] for a in agg.values())
print(f"# {len(rows)} launches, {tot / 1e6:.2f} ms summed (cold-cache, serialised: compare shares, not absolutes)")
for n, (ns, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{ns / 1e6:8.3f} ms  {100 * ns / tot:5.1f}%  x{c:<4d} {n}")
