#!/bin/bash
# Row f3, fused exchange against the NCCL exchange on N GPUs (gpurun --gpus N).
N=${1:-2}
mkdir -p gpurun_out; out=gpurun_out/r2_stripe_fused_${N}gpu.jsonl; : > $out; err=gpurun_out/stripe_fused_${N}.err; : > $err
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 120 $TR tools/stripe_parity.py --workload replica200k --fused >> $out 2>> $err || { echo "fused parity FAILED"; tail -15 $err; }
for wl in hd2m replica200k; do
  for mode in "" "--stripe-fused"; do
    timeout 300 $TR bench.py --gpus $N --split tiles --workload $wl --steps 5 --warmup 2 --iters 50 $mode >> $out 2>> $err || tail -5 $err
  done
done
python - <<PY
import json
for l in open("$out"):
    if not l.startswith("{"): continue
    d = json.loads(l)
    if "value" in d: print(d["n_gpus"], d["config"]["workload"][:28], d["config"].get("exchange"), d["value"], d["e2e"]["value"], d["config"]["us_per_iteration"], d["pose_errors"]["eT_m"])
    else: print(d)
PY
