#!/bin/bash
# BASELINE.json configs 3-5 once each (not the headline bench line): writes gpurun_out/configs.jsonl
mkdir -p gpurun_out; out=gpurun_out/configs.jsonl; : > $out
timeout 600 python bench.py --workload replica200k --cameras 64 --depth-lambda 0.7 --normal-lambda 0.1 --pool 1 --steps 3 --warmup 3 --iters 20 --no-cpu-baseline >> $out 2> gpurun_out/cfg_c64.err || tail -3 gpurun_out/cfg_c64.err
timeout 900 python bench.py --workload hd2m --pool 1 --steps 3 --warmup 3 --iters 50 --no-cpu-baseline >> $out 2> gpurun_out/cfg_hd2m.err || tail -3 gpurun_out/cfg_hd2m.err
timeout 600 python bench.py --workload hd500k --pool 2 --steps 5 --warmup 3 --no-cpu-baseline >> $out 2> gpurun_out/cfg_hd500k.err || tail -3 gpurun_out/cfg_hd500k.err
python - <<'PY'
import json
for l in open("gpurun_out/configs.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:60], "C", d["config"]["hypotheses"], d["value"], d["e2e"]["value"],
          {k: round(v["ms"], 3) for k, v in d["roofline"]["stages"].items()})
PY
