"""Times the per-frame setup kernels (rows f1, f2) on a Replica-sized frame: KNN scale initialisation, back-projection,
GT re-render; prints one JSON line.  The CPU figure beside it is scipy's cKDTree on the host (what the reference's
small_gicp round trip amounts to).

    python tools/setup_time.py [--workload replica200k]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import bench
from gsplatloc_b200 import geometry as geo

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="replica200k")
ap.add_argument("--reps", type=int, default=20)
args = ap.parse_args()
fp = bench.make_frames(args.workload, 1, 42)[0]
dev = torch.device("cuda", 0)
pts = torch.tensor(fp.tar_points, device=dev)
depth = torch.tensor(fp.src_depth, device=dev)
K = torch.tensor(fp.K)
T = torch.tensor(fp.src_c2w)


def timed(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {"workload": args.workload, "points": int(pts.shape[0]), "pixels": int(depth.numel())}
out["init_gs_scales_ms"] = round(timed(lambda: geo.init_gs_scales(pts, 5), args.reps), 4)
out["knn_k5_ms"] = round(timed(lambda: geo.knn(pts, 5), args.reps), 4)
out["depth_to_points_ms"] = round(timed(lambda: geo.depth_to_points(depth, K, c2w=T), args.reps), 4)
full = geo.depth_to_points(depth, K, c2w=T)
out["knn_k5_full_image_ms"] = round(timed(lambda: geo.init_gs_scales(full, 5), 5), 4)
out["full_image_points"] = int(full.shape[0])
from scipy.spatial import cKDTree
t0 = time.time()
x = fp.tar_points.astype(np.float64)
cKDTree(x).query(x, k=5, workers=-1)
out["cpu_ckdtree_ms"] = round((time.time() - t0) * 1e3, 1)
out["cpu_cores"] = os.cpu_count()
print(json.dumps(out))
