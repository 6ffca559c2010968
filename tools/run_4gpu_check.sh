#!/bin/bash
# 4-GPU box: the driver's weak-scaling launch of bench.py at 2 and 4 ranks, then the striped frame with both exchanges.
mkdir -p gpurun_out; out=gpurun_out/r2_4gpu.jsonl; : > $out; err=gpurun_out/r2_4gpu.err; : > $err
tr() { echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511"; }
for n in 2 4; do timeout 400 $(tr $n) bench.py --gpus $n --steps 10 --warmup 3 >> $out 2>> $err || tail -5 $err; done
timeout 120 $(tr 4) tools/stripe_parity.py --workload tum77k --fused >> $out 2>> $err || tail -5 $err
for wl in hd2m replica200k; do
  timeout 300 $(tr 4) bench.py --gpus 4 --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 --stripe-fused >> $out 2>> $err || tail -5 $err
done
python - <<PY
import json
for l in open("$out"):
    if not l.startswith("{"): continue
    d = json.loads(l)
    if "value" in d: print(d["n_gpus"], d["config"].get("split"), d["config"]["workload"][:28], d["config"].get("exchange"), d["value"], d["e2e"]["value"], d.get("pose_errors"))
    else: print(d)
PY
