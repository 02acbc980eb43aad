"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel name, launches, total us, share.
    python tools/ncu_launch_summary.py profiles/r01_ncu_launches_v7.csv > profiles/r01_ncu_launches_v7_summary.txt
"""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
name_i, val_i, unit_i = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
acc = OrderedDict()
for r in rows[1:]:
    v = float(r[val_i].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[unit_i], 1.0)
    name = re.sub(r"\(.*", "", r[name_i]).strip()
    n, t = acc.get(name, (0, 0.0))
    acc[name] = (n + 1, t + v)
total = sum(t for _, t in acc.values())
print(f"# {len(rows) - 1} launches, {total:.1f} us in total (cold-cache, serialised under ncu: compare shares, not times)")
print("# kernel, launches, total us, share")
for name, (n, t) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print(f"{name}, {n}, {t:.1f}, {t / total:.3f}")
