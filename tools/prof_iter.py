"""Short driver for ncu: a few eager (un-graphed) iterations of the pose loop on one synthetic frame.

    python tools/prof_iter.py --workload replica200k --iters 3
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
from gsplatloc_b200 import PoseTracker, TrackerConfig

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="replica200k")
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--advance", type=int, default=0, help="graph iterations to run first (moves the pose off the lattice)")
ap.add_argument("--cameras", type=int, default=1)
args = ap.parse_args()
fp = bench.make_frames(args.workload, 1, 42)[0]
dev = torch.device("cuda", 0)
obs = bench.rendered_observation(PoseTracker, fp, dev)
N = fp.tar_points.shape[0]
t = PoseTracker(fp.W, fp.H, args.cameras, TrackerConfig(max_steps=2000, early_stop=False), device=dev)
t.set_map(torch.tensor(fp.tar_points, device=dev), torch.tensor([[1.0, 0, 0, 0]], device=dev).repeat(N, 1),
          torch.tensor(fp.scales, device=dev), torch.ones(N, device=dev))
t.set_observation(obs, fp.K)
t.set_poses(fp.tar_c2w, fp.src_c2w)
if args.advance:
    t.run(args.advance, poll_every=0, use_graph=True)
t.run(args.iters, poll_every=0, use_graph=False)
print(t.counters(), t.states()[0]["loss"])
