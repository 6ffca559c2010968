#!/usr/bin/env python
"""Tile-striped frame over REAL ranks (torchrun, NCCL): every rank's trajectory must be bit-identical to the unstriped
one-GPU tracker run beside it on the same device.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/stripe_parity.py [--workload tum77k]
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
import bench
from gsplatloc_b200 import PoseTracker, TrackerConfig
from gsplatloc_b200.stripe import StripedTracker

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="tum77k")
ap.add_argument("--iters", type=int, default=25)
ap.add_argument("--fused", action="store_true", help="peer stores + barrier kernels instead of NCCL all-reduces")
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
fp = bench.make_frames(args.workload, 1, 42)[0]
N = fp.tar_points.shape[0]


def make():
    t = PoseTracker(fp.W, fp.H, 1, TrackerConfig(early_stop=False), device=dev)
    t.set_map(torch.tensor(fp.tar_points, device=dev), torch.tensor([[1.0, 0, 0, 0]], device=dev).repeat(N, 1),
              torch.tensor(fp.scales, device=dev), torch.ones(N, device=dev))
    t.set_observation(torch.tensor(fp.src_depth, device=dev), fp.K)
    t.set_poses(fp.tar_c2w, fp.src_c2w)
    return t


ref = make()
ref.run(args.iters, poll_every=0, use_graph=True)
rs, rh = ref.states()[0], ref.history(0)
trk = make()
st = StripedTracker(trk, rank, world)
if args.fused:
    st.fuse()
st.iterate(args.iters)
st.finish()
s, h = trk.states()[0], trk.history(0)
same = bool(np.array_equal(s["q"], rs["q"]) and np.array_equal(s["t"], rs["t"]) and np.array_equal(h, rh)
            and np.array_equal(s["g_q"], rs["g_q"]) and np.array_equal(s["g_t"], rs["g_t"]))
d, _ = ref.render()
same_img = bool(torch.equal(st.depth.view(fp.H, fp.W), d[0]))
flag = torch.tensor([int(same), int(same_img)], device=dev)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"workload": args.workload, "world": world, "fused": args.fused, "iters": args.iters, "trajectory_bit_identical_on_every_rank": bool(flag[0]),
                      "summed_image_bit_identical": bool(flag[1]), "eT_m": float(s["eT"]), "own_isects": trk.counters()["n_isects"],
                      "all_isects": ref.counters()["n_isects"]}), flush=True)
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if bool(flag[0]) and bool(flag[1]) else 1)
