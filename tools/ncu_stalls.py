#!/usr/bin/env python
"""Per-instruction stall summary of one kernel from an Nsight Compute report (works without a GPU).

    python tools/ncu_stalls.py gpurun_out/x.ncu-rep raster_bwd [--top 40] [--view sass|cuda]
"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
view = sys.argv[sys.argv.index("--view") + 1] if "--view" in sys.argv else "sass"
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}", "--launch-count", "1"]
if view == "cuda":
    cmd += ["--print-source", "cuda"]
txt = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
st = [h for h in hdr if h.startswith("stall") and "Not Issued" not in h]
def num(r, h):
    try:
        return int(r[ci[h]] or 0)
    except ValueError:
        return 0
tot = sum(num(r, "# Samples") for r in data)
inst = sum(num(r, "Instructions Executed") for r in data)
print(f"{len(data)} lines, samples {tot}, warp instructions {inst}")
agg = {h: sum(num(r, h) for r in data) for h in st}
print("stall totals:", ", ".join(f"{h[6:]} {v} ({100 * v / max(tot, 1):.1f}%)" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:top_n]:
    s = num(r, "# Samples")
    why = {h[6:]: num(r, h) for h in st if num(r, h) > 0.2 * s}
    print(f"{s:7d} {100 * s / max(tot, 1):5.1f}%  exec {num(r, 'Instructions Executed'):9d}  {r[ci['Source']][:100]}  {why}")
