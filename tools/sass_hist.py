"""Opcode histogram of a SASS address range: python tools/sass_hist.py file.sass 0x15e0 0x3170"""
import re, sys, collections
lo, hi = int(sys.argv[2], 16), int(sys.argv[3], 16)
c = collections.Counter()
for line in open(sys.argv[1]):
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if not m: continue
    a = int(m.group(1), 16)
    if not (lo <= a < hi): continue
    toks = m.group(2).split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    c[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith("MUFU") else "")] += 1
tot = sum(c.values())
print("total", tot)
for k, v in c.most_common(): print(f"{v:5d} {k}")
