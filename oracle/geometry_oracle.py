"""TEST INFRASTRUCTURE ONLY — CPU restatement of the per-frame setup functions around GsplatLoc's pose loop
(SURVEY.md §8 rows f1, f2).  Nothing under ``gsplatloc_b200/`` imports this module; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may.

Pinning: ``tests/golden/ref_geometry_small.npz`` holds the outputs of the reference's OWN ``init_gs_scales``,
``transform_points``, ``depth_to_points`` and ``compute_depth_gt`` (``/root/reference/src/my_gsplat/geometry.py``)
run through ``tests/golden/ref_shims.py``; ``tests/test_golden.py`` holds this module to it.  What the fixture cannot
pin: ``small_gicp`` itself is absent (``poetry.lock:3431-3432``), so the KdTree is scipy's cKDTree returning SQUARED
distances as small_gicp's ``batch_knn_search`` does, and ``gsplat.rasterization`` inside
``compute_depth_gt`` is the rasteriser oracle.
"""
from __future__ import annotations

import numpy as np
from scipy.spatial import cKDTree


def knn_sq_dists(points: np.ndarray, k: int) -> np.ndarray:
    """``knn(x, K)`` (src/my_gsplat/utils.py:16-22): small_gicp ``KdTree(PointCloud(x float64)).batch_knn_search(x, k)``
    second return value = squared distances [N, k], ascending, column 0 the point itself."""
    x = np.asarray(points, dtype=np.float64)
    if x.shape[0] == 0:
        return np.zeros((0, k), dtype=np.float64)
    kk = min(k, x.shape[0])
    d, _ = cKDTree(x).query(x, k=kk)
    d = d.reshape(x.shape[0], kk)
    out = np.full((x.shape[0], k), np.inf)
    out[:, :kk] = d * d
    return out


def init_gs_scales(points: np.ndarray, k: int = 5, eps: float = 1e-24, reference_faithful: bool = True) -> np.ndarray:
    """geometry.py:44-66: dist2_avg = (knn(points, k)[:, 1:] ** 2).mean(-1); sqrt(dist2_avg + eps) repeated to [N, 3],
    in float32 like the reference (``to_tensor`` casts the KdTree output to float32 first).
    ``reference_faithful=False``: the values from knn are taken as already squared (RMS neighbour distance)."""
    d = knn_sq_dists(points, k).astype(np.float32)[:, 1:]
    v = d * d if reference_faithful else d
    s = np.sqrt(v.mean(axis=-1, dtype=np.float32) + np.float32(eps)).astype(np.float32)
    return np.repeat(s[:, None], 3, axis=1)


def transform_points(matrix: np.ndarray, points: np.ndarray) -> np.ndarray:
    """geometry.py:23-41: torch.addmm(matrix[:3, 3], points, matrix[:3, :3].t())."""
    m = np.asarray(matrix, dtype=np.float32)
    p = np.asarray(points, dtype=np.float32)
    return (p @ m[:3, :3].T + m[:3, 3]).astype(np.float32)


def depth_to_points(depth: np.ndarray, K: np.ndarray) -> np.ndarray:
    """geometry.py:138-161 over kornia ``depth_to_3d_v2``: depth [H, W] * ((u - cx) / fx, (v - cy) / fy, 1) -> [H*W, 3]."""
    d = np.asarray(depth, dtype=np.float32)
    K = np.asarray(K, dtype=np.float32)
    H, W = d.shape
    v, u = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    xyz = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u)], -1)
    return (d[..., None] * xyz).reshape(-1, 3).astype(np.float32)


def compute_depth_gt(points: np.ndarray, Ks: np.ndarray, c2w: np.ndarray, height: int, width: int,
                     reference_faithful: bool = True) -> np.ndarray:
    """geometry.py:69-135: "ED" render of the cloud with opacity 1, identity quaternions and KNN scales -> [H, W]."""
    import torch

    from . import gsplat_oracle as orc
    pts = torch.tensor(np.asarray(points, dtype=np.float32))
    N = pts.shape[0]
    scales = torch.tensor(init_gs_scales(points, reference_faithful=reference_faithful))
    quats = torch.tensor([1.0, 0.0, 0.0, 0.0]).repeat(N, 1)
    colors = torch.zeros(N, 4, 3)
    c2w_t = torch.tensor(np.asarray(c2w, dtype=np.float32)).reshape(-1, 4, 4)
    Ks_t = torch.tensor(np.asarray(Ks, dtype=np.float32)).reshape(-1, 3, 3)
    r, _, _ = orc.rasterization(means=pts, quats=quats, scales=scales, opacities=torch.ones(N), colors=colors,
                                sh_degree=1, viewmats=torch.linalg.inv(c2w_t), Ks=Ks_t, width=width, height=height,
                                far_plane=1e10, near_plane=1e-2, render_mode="ED", rasterize_mode="classic")
    return r.squeeze(0).squeeze(-1).numpy()


def align_principle_axes(points: np.ndarray) -> np.ndarray:
    """src/data/normalize.py:7-50 in torch on the CPU (float32 like the reference): lower median per axis, unbiased
    covariance, eigh, eigenvectors by descending eigenvalue, determinant flip, [R | -R c]."""
    import torch
    pc = torch.tensor(np.asarray(points, dtype=np.float32))
    centroid = torch.median(pc, dim=0).values
    cov = torch.cov((pc - centroid).t())
    w, v = torch.linalg.eigh(cov)
    v = v[:, w.argsort(descending=True)]
    if torch.det(v) < 0:
        v[:, 0] *= -1
    R = v.t()
    T = torch.eye(4)
    T[:3, :3] = R
    T[:3, 3] = -torch.mv(R, centroid)
    return T.numpy()


def median_cov(points: np.ndarray):
    """The two reductions of align_principle_axes in float64 (numpy): lower median per axis, unbiased covariance."""
    p = np.asarray(points, dtype=np.float64)
    med = np.sort(p, axis=0)[(p.shape[0] - 1) // 2]
    d = p - med
    return med, np.cov(d.T, ddof=1)
