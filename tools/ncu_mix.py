"""Executed opcode mix per warp of one kernel from an ncu report: python tools/ncu_mix.py rep kernel_regex"""
import csv, io, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]; ci = {h: i for i, h in enumerate(hdr)}
def num(r, h):
    try: return int(r[ci[h]] or 0)
    except ValueError: return 0
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[ci["Source"]] != "Source"]
# the first launch only
seen, first = set(), []
for r in data:
    a = r[ci.get("Address", 0)] if "Address" in ci else None
    if a in seen and a is not None: break
    seen.add(a); first.append(r)
c = collections.Counter()
for r in first:
    t = r[ci["Source"]].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    c[op] += num(r, "Instructions Executed")
tot = sum(c.values())
print("warp instructions", tot)
for k, v in c.most_common(28): print(f"{100 * v / tot:5.1f}%  {v:10d}  {k}")
