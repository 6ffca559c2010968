#!/bin/bash
# Row f3 on N GPUs (gpurun --gpus N): NCCL parity of the striped tracker, then the strong-scaling bench of one frame.
N=${1:-2}
mkdir -p gpurun_out; out=gpurun_out/r2_stripe_${N}gpu.jsonl; : > $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/stripe_parity.py --workload replica200k >> $out 2> gpurun_out/stripe_${N}.err || tail -5 gpurun_out/stripe_${N}.err
for wl in hd2m replica200k; do
  [ "$N" = "2" ] && timeout 600 $TR bench.py --gpus $N --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 --no-stripe-graph >> $out 2>> gpurun_out/stripe_${N}.err
  timeout 600 python bench.py --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 >> $out 2>> gpurun_out/stripe_${N}.err
  timeout 600 $TR bench.py --gpus $N --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 >> $out 2>> gpurun_out/stripe_${N}.err || tail -5 gpurun_out/stripe_${N}.err
done
python - <<PY
import json
for l in open("$out"):
    if not l.startswith("{"): continue
    d = json.loads(l)
    if "value" in d: print(d["n_gpus"], d["config"]["workload"][:40], d["value"], d["e2e"]["value"], d["config"]["us_per_iteration"], d["config"].get("cuda_graph"), d["pose_errors"])
    else: print(d)
PY
