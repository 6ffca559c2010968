"""Multi-GPU sharding of the pose-optimisation path: one process per GPU, units are independent
frame pairs (``for i, train_data in enumerate(self.parser)``, gs_trainer_total.py:49-63 — every frame builds a fresh
GSModel and CameraOptModule) or independent pose hypotheses along gsplat's camera axis.  There is NO data-path
collective: each rank optimises its own units; the only exchange is one all-gather of the fixed-size per-frame result
records at the end (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

# one row per optimised frame / hypothesis: what the reference logs at the best step (gs_trainer_total.py:172-180)
RECORD_FIELDS = ("unit", "rank", "iters", "stopped", "loss", "depth_loss", "edge_loss", "normal_loss", "eT", "eR",
                 "best_loss", "best_eT", "best_eR", "best_iter", "q0", "q1", "q2", "q3", "t0", "t1", "t2")
RECORD_WIDTH = len(RECORD_FIELDS)


def frames_of_rank(n_frames: int, rank: int, world: int) -> List[int]:
    """Round-robin: rank r owns frames {f : f mod world == r} (balances the variance of early-stop lengths)."""
    assert 0 <= rank < world
    return list(range(rank, n_frames, world))


def hypotheses_of_rank(n_hyp: int, rank: int, world: int) -> range:
    """Contiguous blocks [r*C/world, (r+1)*C/world) of the camera axis; sizes differ by at most one."""
    assert 0 <= rank < world
    base, rem = divmod(n_hyp, world)
    lo = rank * base + min(rank, rem)
    return range(lo, lo + base + (1 if rank < rem else 0))


def state_to_record(unit: int, rank: int, st: dict) -> List[float]:
    """``PoseTracker.states()[c]`` -> one record row."""
    return [float(unit), float(rank), float(st["iter"]), float(st["stopped"]), float(st["loss"]),
            float(st["depth_loss"]), float(st["edge_loss"]), float(st["normal_loss"]), float(st["eT"]), float(st["eR"]),
            float(st["best_loss"]), float(st["best_eT"]), float(st["best_eR"]), float(st["best_iter"]),
            *[float(x) for x in st["q"]], *[float(x) for x in st["t"]]]


def gather_records(rows: Sequence[Sequence[float]], device: Optional[torch.device] = None, group=None) -> np.ndarray:
    """The one collective of the path: all ranks get every rank's records, sorted by unit id.  Ranks may hold
    different numbers of rows (round-robin remainders), so counts are exchanged first and rows padded."""
    import torch.distributed as dist
    rows = np.asarray(rows, dtype=np.float64).reshape(-1, RECORD_WIDTH)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return rows[np.argsort(rows[:, 0], kind="stable")]
    world = dist.get_world_size(group)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    n = torch.tensor([rows.shape[0]], dtype=torch.int64, device=device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    buf = torch.zeros((cap, RECORD_WIDTH), dtype=torch.float64, device=device)
    if rows.shape[0]:
        buf[:rows.shape[0]] = torch.from_numpy(rows).to(device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = np.concatenate([p[:c].cpu().numpy() for p, c in zip(parts, counts)], axis=0)
    return out[np.argsort(out[:, 0], kind="stable")]


def best_hypothesis(records: np.ndarray, key: str = "loss") -> int:
    """argmin over the gathered hypotheses (S4: 64 hypotheses per frame); ties -> lowest unit id."""
    col = RECORD_FIELDS.index(key)
    vals = np.where(np.isfinite(records[:, col]), records[:, col], np.inf)
    return int(records[int(np.argmin(vals)), 0])


def trajectory_errors(records: np.ndarray, use_best: bool = False) -> dict:
    """ATE / AAE as the reference reports them: RMSE over frames of the per-pair errors (eval/utils.py:113-119,
    logger.py:289-298); ``use_best`` takes the errors at the best-loss step like the trainer's log."""
    eT = records[:, RECORD_FIELDS.index("best_eT" if use_best else "eT")]
    eR = records[:, RECORD_FIELDS.index("best_eR" if use_best else "eR")]
    ok = np.isfinite(eT) & np.isfinite(eR)
    return {"frames": int(records.shape[0]), "ATE_rmse_m": float(np.sqrt(np.mean(eT[ok] ** 2))) if ok.any() else None,
            "AAE_rmse_deg": float(np.sqrt(np.mean(eR[ok] ** 2))) if ok.any() else None}


@dataclass
class SequenceResult:
    records: np.ndarray        # this rank's rows
    iters: int                 # pose iterations this rank executed
    device_ms: float           # CUDA-event time of the tracker loops (setup excluded)


def chained_pose(st: dict, use_best: bool = True):
    """Pose a finished loop hands to the next one: the pose at the best-loss step when early stopping recorded one
    (``update_pose``: "update with best pose after a train loop", model.py:41-48), else the last pose."""
    if use_best and int(st.get("best_iter", -1)) >= 0 and "best_q" in st and "best_t" in st:
        return st["best_q"], st["best_t"]
    return st["q"], st["t"]


def state_iters(st: dict) -> int:
    """Pose iterations a loop actually executed (graph launches after the stop flag are no-ops and not counted)."""
    return int(st["iter"]) + (1 if st.get("stopped") else 0)


def track_frames(frame_ids: Sequence[int], load_frame, tracker, rank: int = 0, n_iters: Optional[int] = None,
                 poll_every: int = 25, chain: Optional[str] = None, chain_best: bool = True) -> SequenceResult:
    """Optimise the given frames one after another on ``tracker`` (a ``PoseTracker`` with C = 1), the way
    ``Runner.train`` walks the parser.  ``load_frame(i)`` returns a dict with ``means, quats, scales, opacities``
    (map of frame i+1 in its own coordinates), ``depth`` (observation of frame i), ``K``, ``c2w_init`` (ground-truth
    pose of frame i+1's predecessor protocol, gs_trainer_total.py:63) and optionally ``c2w_gt``.

    ``chain`` = ``"last"`` or ``"cv"`` turns the walk into odometry over ``frame_ids`` (consecutive frames):
    ``load_frame(i, c2w_map)`` is then called with the TRACKED pose of the previous loop (``None`` for the first frame,
    whose ``c2w_init`` starts the chain) and places the map there; the loop starts at that pose (``"last"``) or at the
    constant-velocity prediction of ``motion.ConstantVelocity`` (``"cv"``, model.py:41-77).  ``chain_best``: the pose
    handed on is the one at the best-loss step (the reference's contract), not the one ``patience`` steps later."""
    if chain not in (None, "last", "cv"):
        raise ValueError(f"chain must be None, 'last' or 'cv', not {chain!r}")
    rows, iters, ms = [], 0, 0.0
    est = cv = None
    for f in frame_ids:
        fr = load_frame(f, est) if chain else load_frame(f)
        init = fr["c2w_init"]
        if chain:
            if est is None:
                from .motion import ConstantVelocity
                est = np.asarray(fr["c2w_init"], dtype=np.float32).reshape(4, 4)
                cv = ConstantVelocity(est)
            init = cv.c2w() if chain == "cv" else est
        tracker.set_map(fr["means"], fr["quats"], fr["scales"], fr["opacities"])
        tracker.set_observation(fr["depth"], fr["K"])
        tracker.set_poses(init, fr.get("c2w_gt"))
        r = tracker.run(n_iters, poll_every=poll_every, use_graph=True)
        ms += r["elapsed_ms"]
        st = tracker.states()[0]
        iters += state_iters(st) if "iter" in st else r["iters"]
        rows.append(state_to_record(f, rank, st))
        if chain:
            from .tracker import quat_t_to_c2w
            est = quat_t_to_c2w(*chained_pose(st, chain_best)).astype(np.float32)
            cv.update_pose(est)                    # the loop's result becomes the current pose ...
            if chain == "cv":
                cv.update_pose()                   # ... and the next loop starts at the prediction
    return SequenceResult(np.asarray(rows, dtype=np.float64).reshape(-1, RECORD_WIDTH), iters, ms)
