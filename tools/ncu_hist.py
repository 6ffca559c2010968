#!/usr/bin/env python
"""Histogram of a kernel's SASS by execution count (how many instructions run how often) from an ncu report.

    python tools/ncu_hist.py gpurun_out/x.ncu-rep raster_bwd
"""
import csv, io, subprocess, sys
from collections import Counter
rep, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]; ci = {h: i for i, h in enumerate(hdr)}
seen, out = set(), []
for r in rows[hi + 1:]:
    if len(r) != len(hdr) or r[ci["Address"]] in seen:
        continue
    seen.add(r[ci["Address"]]); out.append(r)
def n(r, h):
    try: return int(r[ci[h]] or 0)
    except ValueError: return 0
c, s = Counter(), Counter()
for r in out:
    e = n(r, "Instructions Executed"); c[e] += 1; s[e] += n(r, "# Samples")
tot = sum(k * v for k, v in c.items()); ts = sum(s.values())
print(f"{len(out)} SASS lines, {tot} warp instructions, {ts} samples")
for k, v in sorted(c.items(), key=lambda kv: -kv[0] * kv[1])[:12]:
    print(f"exec {k:9d} x {v:4d} lines = {k * v:11d} ({100 * k * v / tot:5.1f}% of instructions, {100 * s[k] / max(ts, 1):5.1f}% of samples)")
if "--dump" in sys.argv:
    lo = int(sys.argv[sys.argv.index("--dump") + 1])
    for i, r in enumerate(out):
        if n(r, "Instructions Executed") >= lo:
            print(i, n(r, "Instructions Executed"), n(r, "# Samples"), r[ci["Source"]][:80])
