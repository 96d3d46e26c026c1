"""Per-kernel share table of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python scripts/ncu_launch_summary.py launches.csv ["command line"] [skip] [count]"""
import csv
import re
import sys

path = sys.argv[1]
cmd = sys.argv[2] if len(sys.argv) > 2 else ""
skip = int(sys.argv[3]) if len(sys.argv) > 3 else 0
count = int(sys.argv[4]) if len(sys.argv) > 4 else 10 ** 9
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        ns = float(r["Metric Value"].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r["Metric Unit"], 1.0)
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").strip()
        rows.append((name, ns))
rows = rows[skip: skip + count]
tot = sum(ns for _, ns in rows)
agg = {}
for name, ns in rows:
    a = agg.setdefault(name, [0.0, 0])
    a[0] += ns
    a[1] += 1
print(f"ncu launch list of `{cmd}` (launches {skip}..{skip + len(rows)}; cold-cache serialised times), total {tot / 1e6:.1f} ms")
for name, (ns, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:40]:
    print(f"{ns / 1e6:10.3f} ms {100 * ns / tot:5.1f}%  x{n:5d}  {name[:100]}")
