"""Rank source lines of an `ncu --page source --csv --print-source cuda,sass` export by executed
warp instructions and by stall samples.  usage: ncu_lines.py export.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = hdr = None; data = []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if r and r[0] == "Line No": hdr = r; iE = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        try: ie = int(r[iE]); smp = int(r[iS])
        except ValueError: continue
        data.append((cur, int(r[0]), ie, smp, r[1][:100]))
tot = sum(x[2] for x in data); stot = sum(x[3] for x in data)
print("warp instructions", tot, "samples", stot)
byfile = {}
for x in data:
    a = byfile.setdefault(x[0], [0, 0]); a[0] += x[2]; a[1] += x[3]
for f, (a, b) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:24s} instr {a / tot * 100:5.1f}%  samples {b / stot * 100:5.1f}%")
for x in sorted(data, key=lambda x: -x[3])[:top]:
    print(f"{x[0]:22s}{x[1]:4d} instr {x[2] / tot * 100:5.1f}% smp {x[3] / stot * 100:5.1f}% {x[4]}")
