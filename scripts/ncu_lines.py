"""Per-source-line executed warp instructions of one kernel from an .ncu-rep captured with --import-source on.
usage: python scripts/ncu_lines.py report.ncu-rep kernel-regex [top]"""
import collections
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 50
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + rx] + (["--launch-skip", sys.argv[4], "--launch-count", "1"] if len(sys.argv) > 4 else []), capture_output=True, text=True).stdout
cur, agg, tot = None, {}, 0
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if not r[0].isdigit():
        continue
    try:
        n, s = int(r[7]), int(r[6])
    except ValueError:
        continue
    key = (cur, int(r[0]))
    old = agg.get(key, (0, 0, ""))
    agg[key] = (old[0] + n, old[1] + s, r[1].strip()[:100])
    tot += n
print("warp instructions", tot)
by = collections.Counter()
for k, v in agg.items():
    by[k[0]] += v[0]
print(dict(by))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0][:20]:20s} {k[1]:4d} {100 * v[0] / tot:5.1f}% samples {v[1]:5d}  {v[2]}")
