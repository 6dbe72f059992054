"""Pivot an `ncu --csv --metrics ...` log into one line per launch: id, kernel, metric values.
usage: python scripts/ncu_csv.py log.csv"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 8]
h = rows[0]
ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
d = {}
for r in rows[1:]:
    d.setdefault((int(r[ii]), r[ki].split("(")[0].replace("<unnamed>::", "")), {})[r[mi]] = float(r[vi].replace(",", ""))
for (i, k), m in sorted(d.items()):
    print(i, k, " ".join(f"{a.split('__')[-1]}={b:.4g}" for a, b in m.items()))
