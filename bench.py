#!/usr/bin/env python
"""Benchmark of the GsplatLoc pose-optimisation hot path (BASELINE.json metric: pose-opt iterations/s at
1200x680 / ~200k Gaussians, plus achieved HBM GB/s of the dominant kernel).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # the CPU restatement of the reference path (oracle/)

A STEP is one frame's pose optimisation: ``--iters`` (default 100, BASELINE.json configs[1]) iterations of
pose -> projection -> tile intersection + sort -> depth raster -> depth+edge loss -> raster backward lifted to the
6-DoF pose -> Adam.  ``value`` = iterations/s with the frame resident in HBM; ``e2e`` = the same through the public
``PoseTracker`` API with HOST (pinned) inputs, the H2D copies of the map and the observation and the D2H read of
the pose state inside the timed region.  Frames are sharded over ranks (one process per GPU); the only collective
is the final gather of per-frame pose errors.  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch

METRIC = "pose-opt iters/sec (raster fwd+bwd+loss) at 1200x680/200k Gaussians; HBM GB/s"
UNIT = "iters/s"
WORKLOADS = {
    # name: (camera dict name, lattice stride, translation cm, rotation deg)
    "replica200k": ("REPLICA", 2, 2.0, 1.0),      # BASELINE.json configs[1]  (S2)
    "replica50k": ("REPLICA", 4, 2.0, 1.0),       # configs[0]               (S1)
    "replica816k": ("REPLICA", 1, 2.0, 1.0),      # every pixel back-projected, as RGBDImage does (src/data/Image.py:29)
    "tum77k": ("TUM", 2, 6.0, 3.0),               # configs[2]               (S3)
    "hd500k": ("HD", 2, 2.0, 1.0),                # towards configs[4]       (S5)
    "hd2m": ("HD", 1, 2.0, 1.0),                  # configs[4]: 1920x1080, 2.07 M Gaussians (S5)
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def probe_gsplat() -> str:
    """BASELINE.md §1 / SURVEY.md §8(d): the '>= 10x gsplat CUDA' target needs gsplat's own CUDA path on this box.
    One guarded import, result recorded in the bench line."""
    try:
        import gsplat  # noqa: F401
        from gsplat import rasterization  # noqa: F401
        return f"importable: gsplat {getattr(gsplat, '__version__', '?')} (not timed here)"
    except Exception as e:   # ModuleNotFoundError on this image: the submodule is empty and there is no wheel
        return f"unavailable: {type(e).__name__}: {e}"


def kernel_counters(workload: str) -> dict:
    """profiles/kernel_counters.json (tools/ncu_counters.py from the committed ncu capture): DRAM bytes and warp
    instructions per launch of every kernel of the iteration."""
    p = os.path.join(ROOT, "profiles", "kernel_counters.json")
    try:
        return json.load(open(p)).get(workload, {})
    except Exception:
        return {}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 - 0.05 or ts > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); smax = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:  # timed region shorter than the sampling period: take whatever we have
            for ts, line in self.rows[-3:]:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[0])); smax = float(f[1])
                except Exception:
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_frames(workload: str, n: int, seed0: int):
    from gsplatloc_b200 import synthetic as syn
    cam_name, stride, tcm, rdeg = WORKLOADS[workload]
    cam = getattr(syn, cam_name)
    return [syn.make_frame_pair(cam, stride=stride, seed=seed0 + i, trans_cm=tcm, rot_deg=rdeg) for i in range(n)]


def algorithmic_bytes(stage: str, C: int, N: int, M: int, P: int, T: int) -> float:
    """SURVEY.md §8(d) per-unit figures (fp32, ED-only, pose-gradient-only) x the units one launch processes."""
    return {
        # proj_fwd + the whole of isect's count / scan / emit (one fused pass into fixed-capacity tile bins)
        "project_bin": C * N * (40 + 28) + C * N * (12 + 4) + 12.0 * T + C * N * (16 + 8) + 12.0 * M,
        "tile_scan": 0.0,                                          # fused away (slot kept for the stage table)
        "bin_emit": 0.0,
        "tile_sort": 8.0 * M + 12.0 * M,                           # one read of the keys, one write of ids (+offsets)
        "raster_fwd": 32.0 * M + 12.0 * P,
        "loss": 20.0 * P,
        "raster_bwd": 32.0 * M + 16.0 * P + 24.0 * C * N,
        "pose_finish": 48.0 * T,
        "pose_begin": 64.0 * C,
    }[stage]


def rendered_observation(trk_cls, fp, device, stream=None):
    """Observation as the reference builds it with normalize=True: a splat render of the source frame's own points
    from its own pose (src/data/dataset.py:362-372, geometry.py:69-135).  Setup, not timed."""
    Ns = fp.src_points.shape[0]
    g = trk_cls(fp.W, fp.H, 1, device=device, stream=stream)
    g.set_map(torch.tensor(fp.src_points, device=device), torch.tensor([[1.0, 0, 0, 0]], device=device).repeat(Ns, 1),
              torch.tensor(fp.src_scales, device=device), torch.ones(Ns, device=device))
    g.set_observation(torch.zeros(fp.H, fp.W, device=device), fp.K)
    g.set_poses(fp.src_c2w)
    g.forward()
    d, _ = g.render()
    g.close()
    return d[0].contiguous()


def run_native(args, rank, world, local_rank):
    import torch.distributed as dist
    from gsplatloc_b200 import PoseTracker, TrackerConfig, _lib, sharding
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: gsplatloc_b200 has no CPU path")
    _lib.lib()  # fail loudly if the extension is not built
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    K, W_, iters = args.steps, args.warmup, args.iters
    n_pool = args.pool
    split_h = args.split == "hypotheses"
    # --split frames (default): rank r owns its own frames.  --split hypotheses (BASELINE.json configs[3]): every rank
    # holds the SAME frames and a contiguous block of the pose hypotheses of each (sharding.hypotheses_of_rank)
    frames = make_frames(args.workload, n_pool, seed0=42 + (0 if split_h else 1000 * rank))
    Ctot = max(1, args.cameras)                                              # pose hypotheses per frame (configs[3])
    my_h = list(sharding.hypotheses_of_rank(Ctot, rank, world)) if split_h else list(range(Ctot))
    Cn = len(my_h)
    if Cn == 0:
        raise RuntimeError(f"--split hypotheses: {Ctot} hypotheses cannot be split over {world} ranks")
    cfg = TrackerConfig(max_steps=args.max_steps, early_stop=False,          # fixed 100 steps for timing (S2)
                        depth_lambda=args.depth_lambda, normal_lambda=args.normal_lambda)

    def start_poses(fp, seed):
        """configs[3] (S4): the start pose perturbed by N(0, (1 cm)^2), N(0, (0.5 deg)^2) per hypothesis."""
        if Ctot == 1:
            return fp.tar_c2w
        from gsplatloc_b200 import synthetic as syn
        rng = np.random.default_rng(seed)
        out = []
        for c in range(Ctot):
            p = fp.tar_c2w.copy()
            if c:
                p[:3, 3] += rng.normal(0, 0.01, 3).astype(np.float32)
                r = np.deg2rad(rng.normal(0, 0.5, 3))
                p[:3, :3] = (p[:3, :3] @ syn._rot_xyz(*r)).astype(np.float32)
            out.append(p)
        return np.stack(out)[my_h]
    host = []
    trackers = []
    stream = torch.cuda.Stream(dev)          # every tracker launches here; the timing events are recorded here
    for fp in frames:
        N = fp.tar_points.shape[0]
        obs = rendered_observation(PoseTracker, fp, dev, stream)
        pin = dict(means=torch.tensor(fp.tar_points).pin_memory(),
                   quats=torch.tensor([[1.0, 0, 0, 0]]).repeat(N, 1).pin_memory(),
                   scales=torch.tensor(fp.scales).pin_memory(), opac=torch.ones(N).pin_memory(),
                   depth=obs.cpu().pin_memory())
        host.append(pin)
        t = PoseTracker(fp.W, fp.H, Cn, cfg, device=dev, stream=stream)
        t.set_map(pin["means"].to(dev), pin["quats"].to(dev), pin["scales"].to(dev), pin["opac"].to(dev))
        t.set_observation(obs, fp.K)
        pin["start"] = start_poses(fp, 7 + len(trackers))
        t.set_poses(pin["start"], fp.src_c2w)
        trackers.append(t)
    fp0 = frames[0]
    N, P = fp0.tar_points.shape[0], fp0.W * fp0.H
    T = ((fp0.W + 15) // 16) * ((fp0.H + 15) // 16)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_resident(s):
        t, fp = trackers[s % n_pool], frames[s % n_pool]
        t.set_poses(host[s % n_pool]["start"], fp.src_c2w)
        t.run(iters, poll_every=0, use_graph=True)
        return t

    def step_e2e(s):
        t, fp, pin = trackers[s % n_pool], frames[s % n_pool], host[s % n_pool]
        t.set_map(pin["means"], pin["quats"], pin["scales"], pin["opac"])        # H2D from pinned memory
        t.set_observation(pin["depth"], fp.K)                                    # H2D
        t.set_poses(pin["start"], fp.src_c2w)
        t.run(iters, poll_every=0, use_graph=True)
        return t.states()[0]                                                     # D2H

    # ---------------------------------------------------------------- device-resident leg (value)
    for s in range(max(W_, n_pool)):      # every tracker of the pool captures its CUDA graph before the timed region
        step_resident(s)
    launches0 = sum(t.counters()["total_launches"] for t in trackers)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.time()
    e0.record(stream)
    for s in range(K):
        step_resident(W_ + s)
    e1.record(stream)
    barrier()
    wall1 = time.time()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(wall0, wall1) if sampler else None
    launches = sum(t.counters()["total_launches"] for t in trackers) - launches0
    if split_h:   # one row per hypothesis of the first frame: unit = global hypothesis index
        records = [sharding.state_to_record(my_h[c], rank, st) for c, st in enumerate(trackers[0].states())]
    else:
        records = [sharding.state_to_record(i * world + rank, rank, t.states()[0]) for i, t in enumerate(trackers)]
    M = int(np.mean([t.counters()["n_isects"] for t in trackers]))

    # ---------------------------------------------------------------- end-to-end leg (host buffers)
    for s in range(max(min(W_, 2), 1)):
        step_e2e(s)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(stream)
    for s in range(K):
        step_e2e(W_ + s)
    e3.record(stream)
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    h2d = int(sum(v.numel() * 4 for v in host[0].values() if torch.is_tensor(v))) + 16 * 4 * (Cn + 1) + 9 * 4
    d2h = int(__import__("ctypes").sizeof(_lib.PoseState)) * Cn

    # ---------------------------------------------------------------- per-kernel times (roofline), rank 0 only
    stage_ms = None
    if rank == 0:
        t = trackers[0]
        t.set_poses(host[0]["start"], frames[0].src_c2w)
        t.profile(3)
        stage_ms = t.profile(20)

    # ---------------------------------------------------------------- max over ranks + the one collective
    if world > 1:
        tt = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(tt[0]), float(tt[1])
        lt = torch.tensor([launches], device=dev, dtype=torch.float64)
        dist.all_reduce(lt)
        launches = int(lt[0])
    records = sharding.gather_records(records, dev)        # the path's one collective: final gather of pose errors
    if rank != 0:
        return
    # every hypothesis of every frame advances one iteration per launch set
    total_iters = K * iters * (Ctot if split_h else world * Cn)
    value = total_iters / (ms * 1e-3)
    e2e_v = total_iters / (ms_e2e * 1e-3)
    peak, peak_src = peaks()
    dom = max((k for k in stage_ms if k not in ("pose_begin", "pose_finish", "tile_scan")), key=lambda k: stage_ms[k])
    ab = algorithmic_bytes(dom, Cn, N, M, P * Cn, T * Cn)
    ach = ab / (stage_ms[dom] * 1e-3) / 1e9
    sum_ms = sum(stage_ms.values())
    stages = {k: {"ms": round(v, 5), "share": round(v / sum_ms, 4),
                  "alg_GBps": round(algorithmic_bytes(k, Cn, N, M, P * Cn, T * Cn) / (v * 1e-3) / 1e9, 1) if v > 0 else None}
              for k, v in stage_ms.items()}
    iter_bytes = sum(algorithmic_bytes(k, Cn, N, M, P * Cn, T * Cn) for k in stage_ms)
    out = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W_,
        "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "strong" if split_h else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {fp0.W}x{fp0.H}, N={N} Gaussians, M={M} tile intersections, "
                               f"{iters}-iteration pose optimisation per step (BASELINE.json configs[1])",
                   "iters_per_step": iters, "hypotheses": Ctot, "hypotheses_per_rank": Cn, "split": args.split,
                   "frame_iters_per_s": round(value / Ctot, 2), "depth_lambda": args.depth_lambda,
                   "normal_lambda": args.normal_lambda, "gaussians": N, "isects": M, "pixels": P, "tiles": T,
                   "frames_per_rank": n_pool,
                   "sharding": ("one frame, its pose hypotheses in contiguous blocks over the ranks, final all_gather + argmin"
                                if split_h else "independent frames per rank, final all_gather of pose errors"),
                   "l2": f"inputs larger than L2: {n_pool} frames x ~{round((44*N + 24*P + 12*M + 40*N) / 1e6)} MB of "
                         "buffers cycled between steps; iterations inside a step reuse L2 as the real loop does",
                   "timing": "CUDA events on the launching stream, barrier+synchronize both sides, max over ranks"},
        "e2e": {"value": round(e2e_v, 2), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(ms_e2e / K, 4)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": int(ab), "kernel_ms": round(stage_ms[dom], 5),
                     "iteration": {"algorithmic_bytes": int(iter_bytes), "sum_kernel_ms": round(sum_ms, 5),
                                   "achieved_GBps": round(iter_bytes / (sum_ms * 1e-3) / 1e9, 1),
                                   "frac": round(iter_bytes / (sum_ms * 1e-3) / 1e9 / peak, 4)},
                     "stages": stages},
        "pose_errors": sharding.trajectory_errors(records),
    }
    if split_h:
        out["best_hypothesis"] = {"index": sharding.best_hypothesis(records), "of": int(records.shape[0])}
    kc = kernel_counters(args.workload) if Cn == 1 else {}
    if dom in kc:
        out["roofline"]["traffic"] = kc[dom]["dram_bytes"]
        out["roofline"]["traffic_source"] = f"profiles/kernel_counters.json <- {kc.get('source')}"
    # the roofline that binds the rasterisers is the SM's instruction issue rate, not HBM: warp
    # instructions per launch (ncu, same capture) / live kernel time / (SMs x 4 schedulers x SM clock)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    issue_peak = 148 * 4 * sm_mhz * 1e6
    out["roofline"]["issue"] = {
        "unit": "warp instructions/s", "peak": issue_peak, "sm_mhz": sm_mhz,
        "kernels": {k: {"warp_inst": kc[k]["warp_inst"], "achieved": round(kc[k]["warp_inst"] / (stage_ms[k] * 1e-3), 1),
                        "frac": round(kc[k]["warp_inst"] / (stage_ms[k] * 1e-3) / issue_peak, 4)}
                    for k in stage_ms if k in kc and stage_ms[k] > 0}}
    out["gsplat_cuda"] = probe_gsplat()
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_oracle(args.workload, frames[0], n_timed=args.cpu_iters, warm=1)
    print(json.dumps(out), flush=True)


def run_striped(args, rank, world, local_rank):
    """--split tiles (BASELINE.json configs[4]): ONE frame, its tiles interleaved over the ranks
    (gsplatloc_b200/stripe.py).  A step is the frame's ``--iters`` iterations; two all-reduces per iteration (rendered
    depth image, per-tile pose partials) over NCCL / NVLink, no host synchronisation inside a step.  Strong scaling:
    value = iterations/s of the one frame."""
    import torch.distributed as dist
    from gsplatloc_b200 import PoseTracker, TrackerConfig, _lib
    from gsplatloc_b200.stripe import StripedTracker
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: gsplatloc_b200 has no CPU path")
    _lib.lib()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    K, W_, iters = args.steps, args.warmup, args.iters
    fp = make_frames(args.workload, 1, seed0=42)[0]          # the same frame on every rank
    N = fp.tar_points.shape[0]
    stream = torch.cuda.Stream(dev)
    obs = rendered_observation(PoseTracker, fp, dev, stream)
    pin = dict(means=torch.tensor(fp.tar_points).pin_memory(), quats=torch.tensor([[1.0, 0, 0, 0]]).repeat(N, 1).pin_memory(),
               scales=torch.tensor(fp.scales).pin_memory(), opac=torch.ones(N).pin_memory(), depth=obs.cpu().pin_memory())
    trk = PoseTracker(fp.W, fp.H, 1, TrackerConfig(max_steps=args.max_steps, early_stop=False, depth_lambda=args.depth_lambda),
                      device=dev, stream=stream)
    trk.set_map(pin["means"].to(dev), pin["quats"].to(dev), pin["scales"].to(dev), pin["opac"].to(dev))
    trk.set_observation(obs, fp.K)
    st = StripedTracker(trk, rank, world)
    trk.set_poses(fp.tar_c2w, fp.src_c2w)
    st.iterate(2)                 # sizes the bins, warms NCCL up
    st.finish()
    if args.stripe_fused:
        st.fuse()                 # peer stores from the kernels + two barrier kernels instead of two NCCL all-reduces
        st.iterate(2)
        st.finish()
    if not args.no_stripe_graph:
        st.capture()              # one CUDA graph per iteration: 6 kernels + 2 exchanges

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step(e2e):
        if e2e:
            trk.set_map(pin["means"], pin["quats"], pin["scales"], pin["opac"])
            trk.set_observation(pin["depth"], fp.K)
        trk.set_poses(fp.tar_c2w, fp.src_c2w)
        st.iterate(iters)
        st.finish()
        return trk.states()[0] if e2e else None

    out_ms = []
    for e2e in (False, True):
        for _ in range(max(W_, 1) if not e2e else 1):
            step(e2e)
        sampler = ClockSampler(local_rank) if (rank == 0 and not e2e) else None
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.time()
        e0.record(stream)
        for _ in range(K):
            last = step(e2e)
        e1.record(stream)
        barrier()
        w1 = time.time()
        out_ms.append(e0.elapsed_time(e1))
        if sampler:
            clocks = sampler.stop(w0, w1)
    ms, ms_e2e = out_ms
    M_own = trk.counters()["n_isects"]
    stt = trk.states()[0]
    if world > 1:
        tt = torch.tensor([ms, ms_e2e, float(M_own)], device=dev, dtype=torch.float64)
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm)
        ms, ms_e2e, M_max, M_sum = float(mx[0]), float(mx[1]), int(mx[2]), int(sm[2])
    else:
        M_max = M_sum = M_own
    if rank != 0:
        return
    value = K * iters / (ms * 1e-3)
    h2d = int(sum(v.numel() * 4 for v in pin.values())) + 16 * 4 * 2 + 9 * 4
    print(json.dumps({
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W_,
        "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {fp.W}x{fp.H}, N={N} Gaussians, M={M_sum} tile intersections, ONE frame, "
                               f"tiles interleaved over {world} ranks, {iters} iterations per step (BASELINE.json configs[4])",
                   "split": "tiles", "iters_per_step": iters, "isects_per_rank_max": M_max, "isects_per_rank_mean": M_sum // world,
                   "collectives_per_iteration": 0 if (world == 1 or args.stripe_fused) else 2, "cuda_graph": not args.no_stripe_graph,
                   "exchange": "none" if world == 1 else ("fused peer stores + 2 barrier kernels" if args.stripe_fused else "2 NCCL all-reduces"),
                   "nvlink_bytes_per_iteration_per_rank": int(st.nvlink_bytes_per_iter),
                   "us_per_iteration": round(1e3 * ms / (K * iters), 2),
                   "l2": "one frame resident; iterations reuse L2 as the real loop does",
                   "timing": "CUDA events on the launching stream, barrier+synchronize both sides, max over ranks"},
        "e2e": {"value": round(K * iters / (ms_e2e * 1e-3), 2), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": int(__import__("ctypes").sizeof(_lib.PoseState)), "ms_per_step": round(ms_e2e / K, 4)},
        "gpu_launches": int(K * iters * 6 * world), "clocks": clocks,
        "pose_errors": {"eT_m": float(stt["eT"]), "eR_deg": float(stt["eR"]), "loss": float(stt["loss"])},
        "gsplat_cuda": probe_gsplat()}), flush=True)


def run_dropin(args):
    """--api rasterization: the zero-code-change route of INTEGRATION.md §1 -- GsplatLoc's own loop shape with
    ``from gsplatloc_b200 import rasterization`` in place of gsplat's: CameraOptModule-style quaternion + translation
    parameters in torch, ``rasterization(render_mode="RGB+ED", sh_degree=1)`` exactly as GSModel.forward calls it
    (model.py:195-213), mask + depth + silhouette loss (gs_trainer_total.py:104-152), ``backward()``, two torch Adam
    steps and two ExponentialLR steps per iteration.  One GPU, frame resident; reported beside the tracker's number."""
    from gsplatloc_b200 import _lib, rasterization
    from gsplatloc_b200.losses import compute_depth_loss, compute_silhouette_loss
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: gsplatloc_b200 has no CPU path")
    _lib.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    from gsplatloc_b200 import PoseTracker
    fp = make_frames(args.workload, 1, seed0=42)[0]
    N = fp.tar_points.shape[0]
    means = torch.tensor(fp.tar_points, device=dev)
    quats = torch.tensor([[1.0, 0, 0, 0]], device=dev).repeat(N, 1)
    scales = torch.tensor(fp.scales, device=dev)
    opac_logit = torch.full((N,), 1.0, device=dev)                        # GsConfig.init_opa (model.py:121)
    colors = torch.zeros((N, 4, 3), device=dev)                           # sh0 + shN, sh_degree 1 (model.py:150-160)
    gt = rendered_observation(PoseTracker, fp, dev)[None, :, :, None]     # [1, H, W, 1]
    Ks = torch.tensor(fp.K, device=dev)[None]
    c2w0 = torch.tensor(fp.tar_c2w, device=dev)

    def rotmat(q):   # kornia quaternion (wxyz) -> rotation matrix, normalised (transform.py:50-66)
        w, x, y, z = (q / q.norm())
        return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]).reshape(3, 3)

    def one_frame(n_iters):
        from gsplatloc_b200.motion import rotation_matrix_to_quaternion
        q = torch.tensor(rotation_matrix_to_quaternion(fp.tar_c2w[:3, :3]), device=dev, dtype=torch.float32).requires_grad_(True)
        t = c2w0[:3, 3].clone().requires_grad_(True)
        opts = [torch.optim.Adam([q], lr=5e-4, weight_decay=1e-3), torch.optim.Adam([t], lr=1e-3, weight_decay=1e-3)]
        gamma = 0.2 ** (1.0 / args.max_steps)
        scheds = [torch.optim.lr_scheduler.ExponentialLR(o, gamma=gamma) for o in opts]
        bottom = torch.tensor([[0.0, 0, 0, 1]], device=dev)
        last = None
        for _ in range(n_iters):
            c2w = torch.cat([torch.cat([rotmat(q), t[:, None]], 1), bottom], 0)
            renders, alphas, info = rasterization(means=means, quats=quats, scales=scales, opacities=torch.sigmoid(opac_logit),
                                                  colors=colors, sh_degree=1, viewmats=torch.linalg.inv(c2w)[None], Ks=Ks,
                                                  width=fp.W, height=fp.H, packed=False, absgrad=False, sparse_grad=False,
                                                  far_plane=1e10, near_plane=1e-2, render_mode="RGB+ED",
                                                  rasterize_mode="classic")
            depths = renders[..., 3:4]
            mask = (depths != 0).float()
            dl = compute_depth_loss(depths * mask, gt * mask, loss_type="l1")
            sl = compute_silhouette_loss(depths * mask, gt * mask, loss_type="l1")
            total = dl * args.depth_lambda + sl * (1 - args.depth_lambda)
            total.backward()
            for o in opts:
                o.step(); o.zero_grad(set_to_none=True)
            for sc in scheds:
                sc.step()
            last = total.detach()
        return float(last)   # D2H of the loss: the one host read per frame

    K, W_, iters = args.steps, args.warmup, args.iters
    for _ in range(max(W_, 1)):
        one_frame(min(iters, 10))
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        loss = one_frame(iters)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    value = K * iters / (ms * 1e-3)
    print(json.dumps({"metric": METRIC, "api": "rasterization", "value": round(value, 2), "unit": UNIT, "n_gpus": 1, "steps": K,
                      "warmup": W_, "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": f"{args.workload}: {fp.W}x{fp.H}, N={N} Gaussians; drop-in route: "
                                             "gsplatloc_b200.rasterization(RGB+ED, sh_degree 1) + torch autograd + torch Adam, "
                                             f"{iters} iterations per step", "iters_per_step": iters, "final_loss": loss},
                      "note": "every iteration pays gsplat's own API contract: exact-size isect tensors (one host sync for "
                              "n_isects, as in gsplat), per-call allocations, the general backward (per-Gaussian gradients by "
                              "atomics) and ~40 torch launches; the device-resident tracker is the measured headline"}), flush=True)


def cpu_oracle(workload, fp, n_timed=2, warm=1):
    """The CPU restatement (oracle/, kind 'port': the reference's gsplat/kornia code is absent) on the
    same frame, fp32, all host threads; a bounded sample of ``n_timed`` full iterations (fwd + loss + bwd + Adam)."""
    from oracle import gsplat_oracle as orc
    import gsb_testutil as U
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dt = torch.float32
    means, quats, scales, opac = U.map_tensors(fp, dt)
    times = []
    q0 = torch.tensor(orc.rotmat_to_quat(fp.tar_c2w[:3, :3]), dtype=dt)
    t0 = torch.tensor(fp.tar_c2w[:3, 3], dtype=dt)

    def log(step, row, gq, gt_):
        times.append(time.time())
    t_start = time.time()
    orc.pose_opt_loop(q0, t0, means, quats, scales, opac, torch.tensor(fp.src_depth), torch.tensor(fp.K), fp.W, fp.H,
                      c2w_gt=fp.src_c2w, max_steps=1000, n_steps=warm + n_timed + 1, early_stop=False, log=log)
    # times[i] is stamped after the backward of iteration i; the optimiser step belongs to the next interval
    dts = np.diff(np.array(times))[warm:]
    sec = float(np.mean(dts)) if len(dts) else float(time.time() - t_start)
    return {"value": round(1.0 / sec, 5), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{len(dts)} full iterations of {workload} ({means.shape[0]} Gaussians, {fp.W}x{fp.H}) after "
                      f"{warm} warm-up, torch CPU fp32", "s_per_iter": round(sec, 3)}


def run_reference(args, rank, world):
    """--impl reference: the reference's own implementation cannot be imported or compiled here (gsplat / kornia
    absent), so the arm times the oracle port on the host cores; rank 0 only."""
    if rank != 0:
        return
    frames = make_frames(args.workload, 1, seed0=42)
    steps = max(1, min(args.steps, args.ref_max_steps))
    warm = max(1, min(args.warmup, 1))
    t0 = time.time()
    cb = cpu_oracle(args.workload, frames[0], n_timed=steps, warm=warm)
    fp = frames[0]
    out = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world, "steps": steps,
           "warmup": warm, "ms_per_step": round(1e3 / cb["value"], 2), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"{args.workload}: {fp.W}x{fp.H}, N={fp.tar_points.shape[0]} Gaussians; each step = ONE "
                                  "full pose iteration (bounded sample of the 100-iteration optimisation)",
                      "iters_per_step": 1},
           "cpu_baseline": cb,
           "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "wall_s": round(time.time() - t0, 1)}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="replica200k", choices=sorted(WORKLOADS))
    ap.add_argument("--iters", type=int, default=100, help="pose iterations per step (frame)")
    ap.add_argument("--max-steps", type=int, default=2000, help="LR schedule horizon (GsplatLoc_eval.py --num-iters)")
    ap.add_argument("--pool", type=int, default=4, help="distinct frames cycled per rank")
    ap.add_argument("--cameras", type=int, default=1, help="pose hypotheses per frame (BASELINE.json configs[3]: 64)")
    ap.add_argument("--split", default="frames", choices=["frames", "hypotheses", "tiles"],
                    help="what the ranks shard: independent frames (weak scaling), the pose hypotheses of one frame (strong) "
                         "or the tiles of one frame (strong; two all-reduces per iteration)")
    ap.add_argument("--stripe-fused", action="store_true", help="--split tiles: fused exchange (kernels store into every rank's buffers over NVLink, barrier kernels) instead of NCCL all-reduces")
    ap.add_argument("--no-stripe-graph", action="store_true", help="--split tiles: eager segments instead of one CUDA graph per iteration")
    ap.add_argument("--api", default="tracker", choices=["tracker", "rasterization"],
                    help="tracker: the device-resident loop (headline). rasterization: the drop-in route, one "
                         "gsplatloc_b200.rasterization() + torch autograd + torch Adam per iteration, as model.py:195-213 "
                         "and gs_trainer_total.py:95-152 drive gsplat")
    ap.add_argument("--depth-lambda", type=float, default=0.8)
    ap.add_argument("--normal-lambda", type=float, default=0.0)
    ap.add_argument("--cpu-iters", type=int, default=6, help="timed CPU-oracle iterations (~1.6 s each on 16 cores)")
    ap.add_argument("--ref-max-steps", type=int, default=6)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.api == "rasterization":
        if rank == 0:
            run_dropin(args)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        if args.split == "tiles":
            run_striped(args, rank, world, local_rank)
        else:
            run_native(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
