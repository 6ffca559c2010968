#!/usr/bin/env python
"""Condense an Nsight Compute report into the few numbers the roofline discussion needs.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--out profiles/r1_xxx.md] [--title "..."]

Reads the report with ``ncu -i <rep> --page raw --csv`` (works without a GPU) and prints one markdown table row per
profiled launch: duration, DRAM bytes (the ``traffic`` of bench.py's roofline object), DRAM / L2 / L1 / SM
throughput, issue-slot utilisation, pipe utilisation, occupancy and its limiter, registers, shared memory.
"""
from __future__ import annotations

import argparse
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("smsp__inst_executed.sum", "warp_inst"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
    ("lts__t_sectors.sum", "l2_MB"),   # 32-byte sectors through L2 (lts__t_bytes is not in --set full)
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_static", "smem_static"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__occupancy_limit_registers", "lim_regs"),
    ("launch__occupancy_limit_shared_mem", "lim_smem"),
]


def to_bytes(v: str, unit: str) -> float:
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
    return float(v) * mult


def to_us(v: str, unit: str) -> float:
    mult = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1)
    return float(v) * mult


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--out")
    ap.add_argument("--title", default="")
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    lines = []
    if args.title:
        lines.append(f"# {args.title}\n")
    lines.append(f"source: `{args.report}` (ncu --set full --clock-control none; per-launch values, cold-ish caches, "
                 "serialised)\n")
    names = [lab for key, lab in KEYS if key in col]
    lines.append("| kernel | " + " | ".join(names) + " |")
    lines.append("|---|" + "---|" * len(names))
    for r in data:
        kn = r[col["Kernel Name"]]
        short = kn.split("(")[0].replace("void ", "").replace("gsb::", "")
        cells = []
        for key, lab in KEYS:
            if key not in col:
                continue
            v, u = r[col[key]], units[col[key]]
            try:
                if lab == "time":
                    cells.append(f"{to_us(v, u):.1f} us")
                elif lab.startswith("dram_"):
                    cells.append(f"{to_bytes(v, u) / 1e6:.2f} MB")
                elif lab == "l2_MB":
                    cells.append(f"{float(v) * 32 / 1e6:.1f}")
                elif lab in ("warp_inst", "smem_wavefronts"):
                    cells.append(f"{float(v) / 1e6:.1f} M")
                elif lab.endswith("%") or lab == "thr/inst":
                    cells.append(f"{float(v):.1f}")
                else:
                    cells.append(v.split(".")[0] if v.replace(".", "").isdigit() else v)
            except ValueError:
                cells.append(v)
        lines.append(f"| `{short}` | " + " | ".join(cells) + " |")
    text = "\n".join(lines) + "\n"
    if args.out:
        with open(args.out, "w") as f:
            f.write(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
