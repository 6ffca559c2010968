#!/usr/bin/env python
"""CPU side of the lost-frame analysis: replay frames that tools/track_sequence.py --replay dumped
through the oracle loop (oracle/gsplat_oracle.py pose_opt_loop: the reference's loop restated, fp32, same observation,
same schedule and early-stop settings) and put its loss / eT trajectory beside the CUDA tracker's.

    python tools/replay_oracle.py gpurun_out/replay_frames.npz --iters 150 --out profiles/r2_lost_frames.json
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from oracle import gsplat_oracle as orc

ap = argparse.ArgumentParser()
ap.add_argument("npz")
ap.add_argument("--iters", type=int, default=150)
ap.add_argument("--out", default=None)
ap.add_argument("--frames", default=None)
args = ap.parse_args()
z = np.load(args.npz)
K = torch.tensor(z["K"])
max_steps, patience, warm = (int(x) for x in z["cfg"])
frames = [int(f) for f in (args.frames.split(",") if args.frames else z["frames"])]
H, W = z[f"f{frames[0]}_obs"].shape[-2:]
rows = []
for f in frames:
    means = torch.tensor(z[f"f{f}_means"]); scales = torch.tensor(z[f"f{f}_scales"])
    N = means.shape[0]
    quats = torch.tensor([[1.0, 0, 0, 0]]).repeat(N, 1)
    c2w0, c2w_gt = z[f"f{f}_c2w_init"].astype(np.float64), z[f"f{f}_c2w_gt"].astype(np.float64)
    hist_gpu = z[f"f{f}_history"]
    t0 = time.time()
    n = min(args.iters, hist_gpu.shape[0])
    res = orc.pose_opt_loop(torch.tensor(orc.rotmat_to_quat(c2w0[:3, :3]), dtype=torch.float32),
                            torch.tensor(c2w0[:3, 3], dtype=torch.float32), means, quats, scales, torch.ones(N),
                            torch.tensor(z[f"f{f}_obs"]).reshape(H, W), K, W, H, c2w_gt=c2w_gt, max_steps=max_steps,
                            n_steps=n, early_stop=True, patience=patience, warm=warm)
    h = np.array(res["history"])            # loss, L_d, L_s, eT, eR
    k = h.shape[0]
    row = {"frame": f, "iters_compared": k, "oracle_s": round(time.time() - t0, 1),
           "gpu_final": {"iters": int(z[f"f{f}_best"][4]), "best_iter": int(z[f"f{f}_best"][0]), "best_eT_m": float(z[f"f{f}_best"][1]),
                         "best_eR_deg": float(z[f"f{f}_best"][2])},
           "loss_rel_diff_max": float(np.max(np.abs(hist_gpu[:k, 0] - h[:, 0]) / h[:, 0])),
           "eT_abs_diff_max_m": float(np.max(np.abs(hist_gpu[:k, 4] - h[:, 3]))),
           "trajectory": [{"iter": i, "loss_gpu": float(hist_gpu[i, 0]), "loss_oracle": float(h[i, 0]),
                           "eT_gpu_m": float(hist_gpu[i, 4]), "eT_oracle_m": float(h[i, 3]),
                           "eR_gpu_deg": float(hist_gpu[i, 5]), "eR_oracle_deg": float(h[i, 4])}
                          for i in sorted(set(list(range(0, k, max(1, k // 15))) + [k - 1]))]}
    rows.append(row)
    print(json.dumps({k2: v for k2, v in row.items() if k2 != "trajectory"}), flush=True)
if args.out:
    json.dump(rows, open(args.out, "w"), indent=1)
