#!/bin/bash
# Round 2: the other BASELINE.json configurations and the reference's real per-frame size, one bench line each ->
# gpurun_out/r2_configs.jsonl (copied to profiles/).  One GPU; the multi-GPU hypothesis split is run_hsplit.sh.
mkdir -p gpurun_out; out=gpurun_out/r2_configs.jsonl; : > $out
run() { timeout 900 python bench.py "$@" --no-cpu-baseline >> $out 2>> gpurun_out/r2_configs.err || tail -3 gpurun_out/r2_configs.err; }
run --workload replica816k --pool 2 --steps 5 --warmup 3
run --workload tum77k --steps 10 --warmup 3
run --workload replica200k --cameras 64 --depth-lambda 0.7 --normal-lambda 0.1 --pool 1 --steps 3 --warmup 3 --iters 20
run --workload hd2m --pool 1 --steps 3 --warmup 3 --iters 50
run --workload replica200k --api rasterization --steps 3 --warmup 3 --iters 50
run --workload tum77k --api rasterization --steps 3 --warmup 3 --iters 50
python - <<'PY'
import json
for l in open("gpurun_out/r2_configs.jsonl"):
    d = json.loads(l)
    st = {k: round(v["ms"], 3) for k, v in d.get("roofline", {}).get("stages", {}).items()}
    print(d["config"]["workload"][:70], "|", d.get("api", "tracker"), "C", d["config"].get("hypotheses"), d["value"], d.get("e2e", {}).get("value"), st)
PY
