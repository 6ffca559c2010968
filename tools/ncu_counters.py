#!/usr/bin/env python
"""Per-kernel DRAM traffic and warp-instruction counts of one captured iteration -> profiles/kernel_counters.json, the
file bench.py reads for ``roofline.traffic`` and ``roofline.issue`` (so the numbers in the bench line come from the same
capture as profiles/r2_*_iteration_full.md, not from a constant copied by hand).

    python tools/ncu_counters.py gpurun_out/r2_x_prof.ncu-rep --workload replica200k [--out profiles/kernel_counters.json]
"""
import argparse, csv, io, json, os, subprocess

STAGE = (("project_emit", "project_bin"), ("tile_sort", "tile_sort"), ("raster_fwd", "raster_fwd"), ("loss_", "loss"),
         ("normal_rowsum", "loss"), ("raster_bwd", "raster_bwd"), ("pose_finish", "pose_finish"))
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3,
        "msecond": 1e3}

ap = argparse.ArgumentParser()
ap.add_argument("report")
ap.add_argument("--workload", default="replica200k")
ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "kernel_counters.json"))
args = ap.parse_args()
raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
acc = {}
for r in data:
    name = r[col["Kernel Name"]]
    stage = next((s for k, s in STAGE if k in name), None)
    if stage is None:
        continue
    def val(key):
        return float(r[col[key]]) * UNIT.get(units[col[key]], 1.0)
    a = acc.setdefault(stage, {"n": 0, "dram_bytes": 0.0, "warp_inst": 0.0, "ncu_time_us": 0.0, "kernel": name.split("(")[0]})
    a["n"] += 1
    a["dram_bytes"] += val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    a["warp_inst"] += val("smsp__inst_executed.sum")
    a["ncu_time_us"] += val("gpu__time_duration.sum")
out = json.load(open(args.out)) if os.path.exists(args.out) else {}
out["_comment"] = ("per launch, mean over the captured launches: dram__bytes_read.sum + dram__bytes_write.sum, "
                   "smsp__inst_executed.sum, gpu__time_duration.sum from ncu --set full --clock-control none; written by "
                   "tools/ncu_counters.py from the report named in 'source'")
out[args.workload] = {"source": os.path.basename(args.report),
                      **{s: {"kernel": a["kernel"], "dram_bytes": int(a["dram_bytes"] / a["n"]),
                             "warp_inst": int(a["warp_inst"] / a["n"]), "ncu_time_us": round(a["ncu_time_us"] / a["n"], 2)}
                         for s, a in acc.items()}}
json.dump(out, open(args.out, "w"), indent=1)
print(json.dumps(out[args.workload], indent=1))
