"""Device-resident pose optimisation: the inner loop of ``Runner.train``
(``src/my_gsplat/gs_trainer_total.py:78-267``) for one frame pair and ``C`` pose hypotheses, run as six
hand-written kernels per iteration inside a CUDA graph (``csrc/tracker.cu``).  Host code only feeds buffers, polls
the early-stop flag and reads the result back.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib


@dataclass
class TrackerConfig:
    """Numeric defaults of the reference: ``CameraConfig`` (model.py:18-23), ``OptimizationConfig``
    (data/base.py:22-43), ``GsConfig`` (model.py:119-133), scheduler gamma (gs_trainer_total.py:65-72)."""
    max_steps: int = 1000
    trans_lr: float = 1e-3
    quat_lr: float = 5e-4
    quat_opt_reg: float = 1e-3
    trans_opt_reg: float = 1e-3
    depth_lambda: float = 0.8
    normal_lambda: float = 0.0
    early_stop: bool = True
    patience: int = 200
    warmup: int = 100
    near_plane: float = 1e-2
    far_plane: float = 1e10
    eps2d: float = 0.3
    radius_clip: float = 0.0
    normal_mode: str = "reference"     # "reference": cosine over dim=1 as loss.py:93 computes it; "pixel": per pixel

    def to_c(self) -> _lib.TrackerConfig:
        c = _lib.TrackerConfig()
        _lib.lib().gsb_tracker_default_config(ctypes.byref(c), int(self.max_steps))
        c.lr_quat, c.lr_trans = self.quat_lr, self.trans_lr
        c.wd_quat, c.wd_trans = self.quat_opt_reg, self.trans_opt_reg
        c.depth_lambda, c.normal_lambda = self.depth_lambda, self.normal_lambda
        c.early_stop, c.patience, c.warmup = int(self.early_stop), int(self.patience), int(self.warmup)
        c.near_plane, c.far_plane, c.eps2d, c.radius_clip = self.near_plane, self.far_plane, self.eps2d, self.radius_clip
        c.normal_mode = {"off": 0, "pixel": 1, "reference": 2}[self.normal_mode]
        return c


def _state_dict(s: _lib.PoseState) -> dict:
    out = {}
    for name, _ in s._fields_:
        v = getattr(s, name)
        out[name] = np.array(v[:], dtype=np.float32) if hasattr(v, "__len__") else v
    return out


def quat_t_to_c2w(q: np.ndarray, t: np.ndarray) -> np.ndarray:
    """``CameraOptModule_quat_tans.forward`` on the host (model.py:79-82), float64."""
    q = np.asarray(q, dtype=np.float64)
    w, x, y, z = q / max(np.linalg.norm(q), 1e-12)
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    c2w = np.eye(4)
    c2w[:3, :3], c2w[:3, 3] = R, np.asarray(t, dtype=np.float64)
    return c2w


class PoseTracker:
    def __init__(self, width: int, height: int, n_cameras: int = 1, config: Optional[TrackerConfig] = None,
                 device: str | torch.device = "cuda:0", stream: Optional["torch.cuda.Stream"] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("PoseTracker needs a CUDA device (gsplatloc_b200 has no CPU path)")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.W, self.H, self.C = int(width), int(height), int(n_cameras)
        self.config = config or TrackerConfig()
        self._L = _lib.lib()
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            # every kernel of the tracker runs on ONE stream: the caller's, or torch's current stream when that is
            # not the legacy default stream (graphs cannot be captured there), or a new one.  Record CUDA events on
            # ``tracker.stream`` to time it.
            if stream is None:
                cur = torch.cuda.current_stream(self.device)
                stream = cur if cur.cuda_stream != 0 else torch.cuda.Stream(self.device)
            self.stream = stream
            _lib.check(self._L.gsb_tracker_create(ctypes.byref(self._h), self.C, self.W, self.H,
                                                  self.stream.cuda_stream), "tracker_create")
        self._keep = {}
        self.set_config(self.config)

    def _order_after_producer(self, *tensors):
        """CUDA inputs were produced on torch's current stream: make the tracker's stream wait for it."""
        if any(isinstance(t, torch.Tensor) and t.is_cuda for t in tensors):
            self.stream.wait_stream(torch.cuda.current_stream(self.device))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.gsb_tracker_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ inputs
    def set_config(self, config: TrackerConfig):
        self.config = config
        c = config.to_c()
        _lib.check(self._L.gsb_tracker_set_config(self._h, ctypes.byref(c)), "tracker_set_config")

    @staticmethod
    def _f32(x, shape=None) -> torch.Tensor:
        t = torch.as_tensor(x)
        t = t.detach().to(torch.float32).contiguous()
        if shape is not None:
            t = t.reshape(shape)
        return t

    def set_map(self, means, quats, scales, opacities):
        """Gaussians of the frame (``GSModel``, model.py:136-175). CUDA tensors are used in place (D2D copy);
        CPU tensors / numpy arrays are copied H2D (pin them for an asynchronous copy)."""
        means = self._f32(means)
        N = means.shape[0]
        quats, scales, opacities = self._f32(quats, (N, 4)), self._f32(scales, (N, 3)), self._f32(opacities, (N,))
        on_host = not means.is_cuda
        assert all(t.is_cuda != on_host for t in (quats, scales, opacities)), "map tensors must share a device"
        self._keep["map"] = (means, quats, scales, opacities)
        self._order_after_producer(means, quats, scales, opacities)
        _lib.check(self._L.gsb_tracker_set_map(self._h, N, means.data_ptr(), quats.data_ptr(), scales.data_ptr(),
                                               opacities.data_ptr(), int(on_host)), "tracker_set_map")
        self.N = N

    def set_observation(self, depth, K):
        depth = self._f32(depth, (self.H, self.W))
        K = np.ascontiguousarray(np.asarray(torch.as_tensor(K).detach().cpu(), dtype=np.float32).reshape(3, 3))
        self._keep["obs"] = (depth, K)
        self._order_after_producer(depth)
        _lib.check(self._L.gsb_tracker_set_observation(self._h, depth.data_ptr(), int(not depth.is_cuda),
                                                       K.ctypes.data), "tracker_set_observation")

    def set_poses(self, c2w_init, c2w_gt=None):
        init = np.ascontiguousarray(np.asarray(torch.as_tensor(c2w_init).detach().cpu(), dtype=np.float32))
        init = np.broadcast_to(init.reshape(-1, 4, 4), (self.C, 4, 4)).copy()
        gt_p = None
        if c2w_gt is not None:
            gt = np.ascontiguousarray(np.asarray(torch.as_tensor(c2w_gt).detach().cpu(), dtype=np.float32))
            gt = np.broadcast_to(gt.reshape(-1, 4, 4), (self.C, 4, 4)).copy()
            gt_p = gt.ctypes.data
        _lib.check(self._L.gsb_tracker_set_poses(self._h, init.ctypes.data, gt_p), "tracker_set_poses")

    # ------------------------------------------------------------------ run
    def run(self, n_iters: Optional[int] = None, poll_every: int = 25, use_graph: bool = True) -> dict:
        n_iters = int(self.config.max_steps if n_iters is None else n_iters)
        done, ms = ctypes.c_int(0), ctypes.c_float(0.0)
        _lib.check(self._L.gsb_tracker_run(self._h, n_iters, int(poll_every), int(use_graph), ctypes.byref(done),
                                           ctypes.byref(ms)), "tracker_run")
        return {"iters": done.value, "elapsed_ms": ms.value}

    def forward(self):
        _lib.check(self._L.gsb_tracker_forward(self._h), "tracker_forward")

    STAGES = ("pose_begin", "project_bin", "tile_scan", "bin_emit", "tile_sort", "raster_fwd", "loss", "raster_bwd",
              "pose_finish")

    def profile(self, n_iters: int = 10) -> dict:
        """Mean device time (ms) of every kernel of an iteration, CUDA events on the tracker's stream."""
        buf = (ctypes.c_float * len(self.STAGES))()
        _lib.check(self._L.gsb_tracker_profile(self._h, int(n_iters), ctypes.cast(buf, ctypes.c_void_p)),
                   "tracker_profile")
        return dict(zip(self.STAGES, [float(x) for x in buf]))

    # ------------------------------------------------------------------ outputs
    def states(self) -> list:
        arr = (_lib.PoseState * self.C)()
        _lib.check(self._L.gsb_tracker_get_states(self._h, ctypes.cast(arr, ctypes.c_void_p)), "tracker_get_states")
        return [_state_dict(s) for s in arr]

    def poses(self) -> np.ndarray:
        return np.stack([quat_t_to_c2w(s["q"], s["t"]) for s in self.states()])

    def render(self):
        d = torch.empty((self.C, self.H, self.W), dtype=torch.float32, device=self.device)
        a = torch.empty_like(d)
        self._order_after_producer(d)
        _lib.check(self._L.gsb_tracker_get_render(self._h, d.data_ptr(), a.data_ptr(), 0), "tracker_get_render")
        return d, a

    def history(self, camera: int = 0, max_rows: int = 4096) -> np.ndarray:
        buf = np.zeros((max_rows, 8), dtype=np.float32)
        rows = self._L.gsb_tracker_get_history(self._h, camera, buf.ctypes.data, max_rows)
        if rows < 0:
            _lib.check(1, "tracker_get_history")
        return buf[:rows]

    def counters(self) -> dict:
        m, cap, tot, lpi = _lib.c_i64(0), _lib.c_i64(0), _lib.c_i64(0), ctypes.c_int(0)
        _lib.check(self._L.gsb_tracker_get_counters(self._h, ctypes.byref(m), ctypes.byref(cap), ctypes.byref(lpi),
                                                    ctypes.byref(tot)), "tracker_get_counters")
        return {"n_isects": m.value, "capacity": cap.value, "launches_per_iter": lpi.value,
                "total_launches": tot.value}

    def debug_limit_bins(self, tile_cap: int):
        """Test aid: cap the per-tile bins so that the next ``run`` overflows (``RuntimeError``, code 3)."""
        _lib.check(self._L.gsb_tracker_debug_limit_bins(self._h, int(tile_cap)), "tracker_debug_limit_bins")

    def debug_loss_grads(self):
        """d loss / d D and d loss / d alpha [C,H,W] of the last iteration (copies), for parity tests."""
        vd = torch.empty((self.C, self.H, self.W), dtype=torch.float32, device=self.device)
        va = torch.empty_like(vd)
        self._order_after_producer(vd)
        _lib.check(self._L.gsb_tracker_debug_loss_grads(self._h, vd.data_ptr(), va.data_ptr()), "tracker_debug_loss_grads")
        return vd, va

    def debug_tensors(self):
        """Sorted flatten_ids [M], tile offsets [C*T+1] and projected records [C*N,8] of the last forward
        (copies), for parity tests."""
        M = self.counters()["n_isects"]
        tw, th = (self.W + 15) // 16, (self.H + 15) // 16
        fl = torch.empty(M, dtype=torch.int32, device=self.device)
        off = torch.empty(self.C * tw * th + 1, dtype=torch.int32, device=self.device)
        rec = torch.empty((self.C * self.N, 8), dtype=torch.float32, device=self.device)
        self._order_after_producer(rec)
        _lib.check(self._L.gsb_tracker_debug_copy(self._h, fl.data_ptr(), off.data_ptr(), rec.data_ptr()),
                   "tracker_debug_copy")
        return fl, off, rec
