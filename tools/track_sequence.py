#!/usr/bin/env python
"""BASELINE.json configs[2]: a TUM-shaped (640x480) synthetic sequence tracked frame by frame the way
``Runner.train`` walks the parser (gs_trainer_total.py:49-63: a fresh map and pose per frame pair, initialised at the
previous frame's ground-truth pose), frames sharded round-robin over the ranks, one gather of pose errors at the end.

Per frame, everything after the depth images (the dataset's job) runs on the GPU:
    depth_to_points + transform_points (fused)  ->  lattice subsample  ->  init_gs_scales (grid KNN)
    ->  PoseTracker.set_map / set_observation / set_poses  ->  pose loop with the reference's early stop
The depth images are ray cast on the host BEFORE the timed region (they stand for the PNGs of the dataset).

    python tools/track_sequence.py --frames 64                                        # one GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/track_sequence.py --frames 64

Rank 0 prints one JSON line: frames/s and iterations/s of the whole job (max over ranks of the wall time of the
timed region), the share of the per-frame setup, ATE / AAE over all frames.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--camera", default="TUM", choices=["TUM", "REPLICA"])
    ap.add_argument("--stride", type=int, default=2)
    ap.add_argument("--max-steps", type=int, default=1000, help="GsplatLoc_eval.py --num-iters")
    ap.add_argument("--patience", type=int, default=200)
    ap.add_argument("--trans-cm", type=float, default=6.0)
    ap.add_argument("--rot-deg", type=float, default=3.0)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--obs", default="rendered", choices=["rendered", "raw"],
                    help="observation: compute_depth_gt re-render of the source frame (dataset.py:362-372) or raw depth")
    ap.add_argument("--chain", default="off", choices=["off", "last", "cv"],
                    help="off: every loop starts at the previous frame's GROUND-TRUTH pose, as the shipped trainer does "
                         "(gs_trainer_total.py:49-63), frames round-robin over ranks.  last / cv: odometry -- the map is "
                         "placed at the TRACKED pose of the previous loop and the loop starts there (last) or at the "
                         "constant-velocity prediction (cv: model.py:41-77); each rank chains a contiguous block")
    ap.add_argument("--dump-records", default=None, help="rank 0 writes the per-frame result rows (sharding.RECORD_FIELDS) here (.npy)")
    ap.add_argument("--replay", default=None,
                    help="comma-separated frame ids: after the walk, rank 0 tracks these frames again and writes their inputs "
                         "(map, observation, poses) and per-iteration history to --replay-out (.npz) for tools/replay_oracle.py")
    ap.add_argument("--replay-out", default="gpurun_out/replay_frames.npz")
    ap.add_argument("--replay-only", action="store_true", help="skip the walk: ray cast and track the --replay frames only")
    args = ap.parse_args()

    import torch.distributed as dist
    from gsplatloc_b200 import PoseTracker, TrackerConfig, geometry as geo, motion, sharding, synthetic as syn
    from gsplatloc_b200.tracker import quat_t_to_c2w

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise RuntimeError("track_sequence.py needs a CUDA device: gsplatloc_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    cam = getattr(syn, args.camera)
    chained = args.chain != "off"
    mine = (list(sharding.hypotheses_of_rank(args.frames, rank, world)) if chained
            else sharding.frames_of_rank(args.frames, rank, world))
    if args.replay and args.replay_only:
        mine = [int(x) for x in args.replay.split(",")] if rank == 0 else []
    # ---- the "dataset": every rank ray casts the whole walk (cheap, deterministic) and keeps its own frames
    rng = np.random.default_rng(args.seed)
    scene = syn.make_scene(rng)
    pose = syn.make_pose(rng)
    W, H = cam["W"], cam["H"]
    K = np.array([[cam["fx"], 0, cam["cx"]], [0, cam["fy"], cam["cy"]], [0, 0, 1]], dtype=np.float64)
    frames = {}
    for i in range(args.frames):
        nxt = syn.perturb_pose(pose, np.random.default_rng(args.seed * 100003 + i), args.trans_cm, args.rot_deg)
        if i in mine:
            frames[i] = dict(depth_tar=torch.tensor(syn.raycast_depth(scene, pose, K, W, H)).pin_memory(),
                             depth_src=torch.tensor(syn.raycast_depth(scene, nxt, K, W, H)).pin_memory(),
                             c2w_tar=pose.astype(np.float32), c2w_src=nxt.astype(np.float32))
        pose = nxt.copy()
        pose[:3, 3] = np.clip(pose[:3, 3], [-1.0, -0.5, -1.0], [1.0, 0.5, 1.0])
    K32 = K.astype(np.float32)
    Kt = torch.tensor(K32)

    cfg = TrackerConfig(max_steps=args.max_steps, early_stop=True, patience=args.patience)
    trk = PoseTracker(W, H, 1, cfg, device=dev)
    s = args.stride

    def setup(fr, c2w_map):
        """depth image of the target frame -> map of the pose loop (geometry.py:138-161,23-41,44-66; model.py:136-178)."""
        pts = geo.depth_to_points(fr["depth_tar"].to(dev, non_blocking=True), Kt, c2w=torch.tensor(c2w_map))
        pts = pts.view(H, W, 3)[s // 2::s, s // 2::s].reshape(-1, 3).contiguous()
        scales = geo.init_gs_scales(pts, 5, reference_faithful=False)
        n = pts.shape[0]
        quats = torch.zeros((n, 4), device=dev)
        quats[:, 0] = 1.0
        return pts, quats, scales, torch.ones(n, device=dev)

    eye = torch.eye(4)[None].to(dev)
    Ks_dev = Kt[None].to(dev)

    def observation(fr):
        """dataset.py:362-372 (Parser with normalize=True): the observed depth the loop compares against is the "ED"
        splat render of the source frame's OWN back-projected lattice from its OWN viewpoint (compute_depth_gt), so
        that map and observation carry the same splatting bias; the raw sensor depth is the other choice."""
        d = fr["depth_src"].to(dev, non_blocking=True)
        if args.obs == "raw":
            return d
        cam_pts = geo.depth_to_points(d, Kt).view(H, W, 3)[s // 2::s, s // 2::s].reshape(-1, 3).contiguous()
        return geo.compute_depth_gt(cam_pts, None, Ks_dev, eye, H, W, reference_faithful=False).contiguous()

    def track(fr, c2w_map=None, c2w_init=None):
        t0 = time.perf_counter()
        c2w_map = fr["c2w_tar"] if c2w_map is None else c2w_map
        m = setup(fr, c2w_map)
        trk.set_map(*m)
        trk.set_observation(observation(fr), K32)
        trk.set_poses(c2w_map if c2w_init is None else c2w_init, fr["c2w_src"])
        torch.cuda.synchronize(dev)
        t1 = time.perf_counter()
        r = trk.run(None, poll_every=25, use_graph=True)
        st = trk.states()[0]
        return st, r, t1 - t0, time.perf_counter() - t1

    if mine:
        track(frames[mine[0]])                     # warm-up: graph capture, allocator pools, module load
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    w0 = time.perf_counter()
    rows, iters, t_setup, t_loop = [], 0, 0.0, 0.0
    est = cv = None
    for f in mine:
        if chained:
            if est is None:                            # a block starts at its first frame's ground-truth pose
                est = frames[f]["c2w_tar"]
                cv = motion.ConstantVelocity(est)
            st, r, ts, tl = track(frames[f], est, cv.c2w() if args.chain == "cv" else est)
            est = quat_t_to_c2w(*sharding.chained_pose(st)).astype(np.float32)   # the best-loss pose (model.py:41-48)
            cv.update_pose(est)                        # the loop's result becomes the current pose ...
            if args.chain == "cv":
                cv.update_pose()                       # ... and the next loop starts at the prediction
        else:
            st, r, ts, tl = track(frames[f])
        rows.append(sharding.state_to_record(f, rank, st))
        iters += sharding.state_iters(st); t_setup += ts; t_loop += tl
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - w0
    agg = torch.tensor([wall, float(iters), t_setup, t_loop], device=dev, dtype=torch.float64)
    if world > 1:
        mx = agg.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = agg.clone(); dist.all_reduce(sm)
        wall, iters, t_setup, t_loop = float(mx[0]), int(sm[1]), float(sm[2]), float(sm[3])
    records = sharding.gather_records(rows, dev)
    if rank == 0:
        err = sharding.trajectory_errors(records)
        best = sharding.trajectory_errors(records, use_best=True)
        print(json.dumps({
            "workload": f"{args.camera} {W}x{H}, stride {s} lattice ({(H // s) * (W // s)} Gaussians), {args.frames} frames, "
                        f"early stop (patience {args.patience}, max {args.max_steps}), observation {args.obs}, chain {args.chain}",
            "n_gpus": world, "frames_per_s": round(args.frames / wall, 2), "iters_per_s": round(iters / wall, 1),
            "iters_per_frame": round(iters / max(args.frames, 1), 1), "wall_s": round(wall, 3),
            "setup_share": round(t_setup / max(t_setup + t_loop, 1e-9), 4),
            "setup_ms_per_frame": round(1e3 * t_setup / max(args.frames, 1), 3),
            "eT_median_m": float(np.median(records[:, 8])), "eT_max_m": float(records[:, 8].max()),
            "frames_within_1cm": int((records[:, 8] < 1e-2).sum()), "frames_within_1mm": int((records[:, 8] < 1e-3).sum()),
            "iters_min_max": [int(records[:, 2].min()), int(records[:, 2].max())],
            "ATE_rmse_m": err["ATE_rmse_m"], "AAE_rmse_deg": err["AAE_rmse_deg"],
            "ATE_rmse_m_at_best_loss": best["ATE_rmse_m"], "AAE_rmse_deg_at_best_loss": best["AAE_rmse_deg"]}), flush=True)
        if args.dump_records:
            np.save(args.dump_records, records)
        if args.replay:
            out = {}
            for f in [int(x) for x in args.replay.split(",")]:
                fr = frames[f]
                m = setup(fr, fr["c2w_tar"])
                obs = observation(fr)
                trk.set_map(*m)
                trk.set_observation(obs, K32)
                trk.set_poses(fr["c2w_tar"], fr["c2w_src"])
                trk.run(None, poll_every=25, use_graph=True)
                st = trk.states()[0]
                out.update({f"f{f}_means": m[0].cpu().numpy(), f"f{f}_scales": m[2].cpu().numpy(),
                            f"f{f}_obs": obs.cpu().numpy(), f"f{f}_c2w_init": fr["c2w_tar"], f"f{f}_c2w_gt": fr["c2w_src"],
                            f"f{f}_history": trk.history(0), f"f{f}_best": np.array([st["best_iter"], st["best_eT"], st["best_eR"],
                                                                                    st["best_loss"], st["iter"]], dtype=np.float64)})
            out["K"] = K32
            out["frames"] = np.array([int(x) for x in args.replay.split(",")])
            out["cfg"] = np.array([args.max_steps, args.patience, 100], dtype=np.int64)
            np.savez_compressed(args.replay_out, **out)
    trk.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
