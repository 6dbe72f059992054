"""Instruction count and opcode mix per kernel from `cuobjdump -sass lib.so > sass.txt`.
usage: python scripts/sass_count.py sass.txt [kernel-regex]"""
import collections
import re
import sys

name = None
ops = collections.defaultdict(collections.Counter)
for line in open(sys.argv[1]):
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        ops[name][m.group(2).split(".")[0]] += 1
for n, c in ops.items():
    if len(sys.argv) > 2 and not re.search(sys.argv[2], n):
        continue
    short = re.sub(r"^_ZN?\d*_GLOBAL__N__\w+?_cu_[0-9a-f]+\d+", "", n)[:44]
    print(sum(c.values()), short, c.most_common(8))
