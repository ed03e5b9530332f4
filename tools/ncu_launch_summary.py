"""Summarise an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...):
   python tools/ncu_launch_summary.py X.csv > profiles/..._summary.txt"""
import csv, re, sys
from collections import defaultdict
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]; K = h.index("Kernel Name"); V = h.index("Metric Value"); U = h.index("Metric Unit")
tot = defaultdict(float); cnt = defaultdict(int)
for r in rows[1:]:
    name = re.sub(r"^void ", "", r[K]); name = re.sub(r"\(.*$", "", name)
    t = float(r[V].replace(",", "")) * ({"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[U], 1e-3))
    tot[name] += t; cnt[name] += 1
allt = sum(tot.values())
print("%-90s %8s %12s %8s %8s" % ("kernel", "launches", "total_us", "avg_us", "share"))
for k in sorted(tot, key=lambda k: -tot[k]):
    print("%-90s %8d %12.1f %8.2f %7.1f%%" % (k[:90], cnt[k], tot[k], tot[k] / cnt[k], 100 * tot[k] / allt))
