#!/bin/bash
# One call on an 8-GPU box: row f3 (tile-striped frame) at 8 ranks, and BASELINE.json configs[3] (64 pose hypotheses of
# one frame split over 2 / 4 / 8 ranks).  Lines go to gpurun_out/r2_8gpu.jsonl.
mkdir -p gpurun_out; out=gpurun_out/r2_8gpu.jsonl; : > $out; err=gpurun_out/r2_8gpu.err; : > $err
tr() { echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511"; }
timeout 300 $(tr 8) tools/stripe_parity.py --workload replica200k >> $out 2>> $err || tail -5 $err
for wl in hd2m replica200k; do
  timeout 600 $(tr 8) bench.py --gpus 8 --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 >> $out 2>> $err || tail -5 $err
done
HY="--split hypotheses --cameras 64 --depth-lambda 0.7 --normal-lambda 0.1 --pool 1 --steps 3 --warmup 3 --iters 20 --no-cpu-baseline"
for n in 2 4 8; do
  timeout 600 $(tr $n) bench.py --gpus $n $HY >> $out 2>> $err || tail -5 $err
done
python - <<PY
import json
for l in open("$out"):
    if not l.startswith("{"): continue
    d = json.loads(l)
    if "value" in d: print(d["n_gpus"], d["config"].get("split"), d["config"]["workload"][:34], d["value"], d["e2e"]["value"], d["config"].get("us_per_iteration"), d["config"].get("frame_iters_per_s"), d.get("best_hypothesis"))
    else: print(d)
PY
