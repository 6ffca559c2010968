"""Import shims that let GsplatLoc's own Python modules (``/root/reference/src/...``) be imported in the build
container, where none of their third-party dependencies is installed.  Used ONLY by
``tests/golden/make_golden.py`` to generate the committed fixtures; nothing in the product or in the GPU tests
imports this file, and ``/root/reference`` is never read at test time.

What is shimmed, and how faithfully:

* ``kornia`` 0.7.2 (poetry.lock:1236-1237) — the five functions the hot path calls, restated from the published
  kornia 0.7.2 sources [UPSTREAM-RECALL]: ``filters.sobel`` (loss.py:51-52), ``geometry.depth_to_3d_v2``
  (geometry.py:158), ``geometry.normalize_quaternion`` / ``quaternion_to_rotation_matrix`` /
  ``rotation_matrix_to_quaternion`` (transform.py:65-66,84).
* ``gsplat.rasterization`` (the absent submodule) — bound to the CPU oracle ``oracle.gsplat_oracle.rasterization``.
  Everything that goes through it stays "parity unpinned"; everything around it (losses, normals, pose module,
  Adam, schedulers, metrics) is the reference's own code running unmodified.
* ``small_gicp`` — ``KdTree.batch_knn_search`` on scipy's cKDTree returning SQUARED distances like small_gicp
  (SURVEY.md Appendix C #6).  Setup-time only.
* ``nerfview``, ``viser``, ``torchmetrics``, ``natsort``, ``cv2``, ``wandb``, ``matplotlib``: empty placeholders
  (imported by the reference at module level, never called on this path).
* ``geometry.construct_full_pose`` hard-codes ``device="cuda:0"`` (geometry.py:17): the module's ``torch`` name is
  rebound to a proxy whose ``eye`` drops the device so the function runs on the CPU; the function body is untouched.
"""
from __future__ import annotations

import importlib
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

REFERENCE_ROOT = "/root/reference"


# ----------------------------------------------------------------------------------------------- kornia 0.7.2
def _sobel(input: torch.Tensor, normalized: bool = True, eps: float = 1e-6) -> torch.Tensor:
    """kornia.filters.sobel: input [B,C,H,W] -> sqrt(gx^2 + gy^2 + eps); spatial_gradient(mode='sobel', order=1):
    replicate padding, correlation with [[-1,0,1],[-2,0,2],[-1,0,1]] and its transpose, each / 8 when normalized."""
    assert input.dim() == 4
    b, c, h, w = input.shape
    kx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]], dtype=input.dtype, device=input.device)
    k = torch.stack([kx, kx.t()])
    if normalized:
        k = k / k.abs().sum(dim=(-2, -1), keepdim=True)
    x = F.pad(input.reshape(b * c, 1, h, w), (1, 1, 1, 1), mode="replicate")
    g = F.conv2d(x, k[:, None]).reshape(b, c, 2, h, w)
    gx, gy = g[:, :, 0], g[:, :, 1]
    return torch.sqrt(gx * gx + gy * gy + eps)


def _depth_to_3d_v2(depth: torch.Tensor, camera_matrix: torch.Tensor, normalize_points: bool = False) -> torch.Tensor:
    """kornia.geometry.depth_to_3d_v2: depth [H,W] (or [B,H,W]), K[3,3] -> points [H,W,3] =
    depth * ((u - cx)/fx, (v - cy)/fy, 1), u = 0..W-1, v = 0..H-1."""
    assert not normalize_points
    H, W = depth.shape[-2:]
    K = camera_matrix
    v, u = torch.meshgrid(torch.arange(H, dtype=depth.dtype, device=depth.device),
                          torch.arange(W, dtype=depth.dtype, device=depth.device), indexing="ij")
    xyz = torch.stack([(u - K[..., 0, 2]) / K[..., 0, 0], (v - K[..., 1, 2]) / K[..., 1, 1], torch.ones_like(u)], -1)
    return depth[..., None] * xyz


def _normalize_quaternion(quaternion: torch.Tensor, eps: float = 1e-12) -> torch.Tensor:
    return F.normalize(quaternion, p=2.0, dim=-1, eps=eps)


def _quaternion_to_rotation_matrix(quaternion: torch.Tensor) -> torch.Tensor:
    """kornia 0.7.2, (w, x, y, z) order; normalises its input first."""
    q = _normalize_quaternion(quaternion)
    w, x, y, z = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    one = torch.ones_like(w)
    m = torch.stack([one - (tyy + tzz), txy - twz, txz + twy,
                     txy + twz, one - (txx + tzz), tyz - twx,
                     txz - twy, tyz + twx, one - (txx + tyy)], -1)
    return m.reshape(quaternion.shape[:-1] + (3, 3))


def _rotation_matrix_to_quaternion(rotation_matrix: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """kornia 0.7.2 trace method with torch.where branches, (w, x, y, z)."""
    m = rotation_matrix.reshape(rotation_matrix.shape[:-2] + (9,))
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.chunk(m, 9, dim=-1)
    trace = m00 + m11 + m22

    def safe_div(n, d):
        tiny = torch.finfo(n.dtype).tiny
        return n / torch.clamp(d, min=tiny)

    def trace_positive():
        sq = torch.sqrt(trace + 1.0 + eps) * 2.0
        return torch.cat((0.25 * sq, safe_div(m21 - m12, sq), safe_div(m02 - m20, sq), safe_div(m10 - m01, sq)), -1)

    def cond_1():
        sq = torch.sqrt(1.0 + m00 - m11 - m22 + eps) * 2.0
        return torch.cat((safe_div(m21 - m12, sq), 0.25 * sq, safe_div(m01 + m10, sq), safe_div(m02 + m20, sq)), -1)

    def cond_2():
        sq = torch.sqrt(1.0 + m11 - m00 - m22 + eps) * 2.0
        return torch.cat((safe_div(m02 - m20, sq), safe_div(m01 + m10, sq), 0.25 * sq, safe_div(m12 + m21, sq)), -1)

    def cond_3():
        sq = torch.sqrt(1.0 + m22 - m00 - m11 + eps) * 2.0
        return torch.cat((safe_div(m10 - m01, sq), safe_div(m02 + m20, sq), safe_div(m12 + m21, sq), 0.25 * sq), -1)

    where_2 = torch.where(m11 > m22, cond_2(), cond_3())
    where_1 = torch.where((m00 > m11) & (m00 > m22), cond_1(), where_2)
    return torch.where(trace > 0.0, trace_positive(), where_1)


# ----------------------------------------------------------------------------------------------- small_gicp
class _PointCloud:
    def __init__(self, pts):
        self.pts = np.asarray(pts, dtype=np.float64)


class _KdTree:
    def __init__(self, pcd, num_threads: int = 1):
        from scipy.spatial import cKDTree
        self.tree = cKDTree(pcd.pts)

    def batch_knn_search(self, queries, k: int, num_threads: int = 1):
        d, i = self.tree.query(np.asarray(queries, dtype=np.float64), k=k)
        return i, d * d      # small_gicp returns squared distances (its own test names them k_sq_dists)


class _TorchOnCpu:
    """``torch`` for geometry.py with ``eye`` ignoring the hard-coded ``device="cuda:0"`` (geometry.py:17)."""

    def __getattr__(self, name):
        return getattr(torch, name)

    @staticmethod
    def eye(*a, **kw):
        kw.pop("device", None)
        return torch.eye(*a, **kw)


def _placeholder(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []   # let ``import name.sub`` resolve against sys.modules
    sys.modules[name] = m
    return m


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return self

    def __getattr__(self, n):
        return _Anything()


def install(oracle_rasterization) -> dict:
    """Install the shims, import the reference's hot-path modules and return them by name."""
    kornia = _placeholder("kornia")
    kornia.filters = _placeholder("kornia.filters", sobel=_sobel)
    kornia.geometry = _placeholder("kornia.geometry", depth_to_3d_v2=_depth_to_3d_v2,
                                   normalize_quaternion=_normalize_quaternion,
                                   quaternion_to_rotation_matrix=_quaternion_to_rotation_matrix,
                                   rotation_matrix_to_quaternion=_rotation_matrix_to_quaternion)
    _placeholder("gsplat", rasterization=oracle_rasterization)
    _placeholder("small_gicp", PointCloud=_PointCloud, KdTree=_KdTree)
    _placeholder("nerfview", Viewer=_Anything, CameraState=_Anything)
    _placeholder("viser", ViserServer=_Anything)
    tm = _placeholder("torchmetrics")
    tm.image = _placeholder("torchmetrics.image", LearnedPerceptualImagePatchSimilarity=_Anything,
                            PeakSignalNoiseRatio=_Anything, StructuralSimilarityIndexMeasure=_Anything)
    _placeholder("natsort", natsorted=sorted)
    for name in ("cv2", "wandb", "matplotlib", "matplotlib.pyplot", "open3d", "tqdm"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                _placeholder(name)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    mods = {}
    importlib.import_module("src.data")      # the reference's own import order (src.data first) avoids its import cycle
    for key, name in (("geometry", "src.my_gsplat.geometry"), ("loss", "src.my_gsplat.loss"),
                      ("transform", "src.my_gsplat.transform"), ("model", "src.my_gsplat.model"),
                      ("eval_utils", "src.eval.utils"), ("base", "src.data.base")):
        mods[key] = importlib.import_module(name)
    mods["geometry"].torch = _TorchOnCpu()
    return mods
