"""Per-frame setup around the pose loop: the functions of the reference's
``src/my_gsplat/geometry.py`` and ``src/my_gsplat/utils.py`` that sit between the dataset and the loop, under the
same names and argument meaning, on hand-written sm_100a kernels behind the C ABI (``csrc/knn.cu``,
``csrc/geometry.cu``).  No CPU path: a non-CUDA tensor raises.

* ``knn``               utils.py:16-22      (small_gicp KdTree on the host, GPU -> CPU -> GPU per frame, in the reference)
* ``init_gs_scales``    geometry.py:44-66
* ``transform_points``  geometry.py:23-41
* ``depth_to_points``   geometry.py:138-161 (kornia ``depth_to_3d_v2``)
* ``compute_depth_gt``  geometry.py:69-135  (ED re-render of the source cloud; dataset.py:362-372)
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
from torch import Tensor

from . import _lib
from .ops import _chk, _p, _stream


def _host16(m) -> "C.Array":
    a = np.ascontiguousarray(m.detach().cpu().numpy() if isinstance(m, Tensor) else m, dtype=np.float32).reshape(-1)
    return (C.c_float * a.size)(*a.tolist())


def knn(x: Tensor, K: int = 4) -> Tensor:
    """``knn(x, K)`` of utils.py:16-22: [N, K] neighbour "distances" of every point to its own cloud, nearest first,
    column 0 being the point itself.  As in the reference the values are what ``small_gicp``'s ``batch_knn_search``
    returns second: SQUARED distances."""
    x = _chk(x.detach(), "x")
    assert x.dim() == 2 and x.shape[1] == 3, x.shape
    out = torch.empty((x.shape[0], K), dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().gsb_knn_sq_dists(x.shape[0], _p(x), K, _p(out), None, _stream()), "knn_sq_dists")
    return out


def init_gs_scales(points: Tensor, k: int = 5, eps: float = 1e-24, reference_faithful: bool = True) -> Tensor:
    """geometry.py:44-66: ``sqrt(mean(knn(points, k)[:, 1:] ** 2) + eps)`` repeated to [N, 3], fused into the KNN
    query kernel.  ``reference_faithful=True`` squares the squared distances again exactly like the reference;
    ``False`` gives the RMS distance to the k-1 nearest neighbours (what the paper describes)."""
    points = _chk(points.detach(), "points")
    assert points.dim() == 2 and points.shape[1] == 3, points.shape
    out = torch.empty((points.shape[0], 3), dtype=torch.float32, device=points.device)
    _lib.check(_lib.lib().gsb_init_gs_scales(points.shape[0], _p(points), k, eps, 0 if reference_faithful else 1,
                                             _p(out), _stream()), "init_gs_scales")
    return out


def transform_points(matrix: Tensor, points: Tensor) -> Tensor:
    """geometry.py:23-41: ``points @ R^T + t`` for a 4x4 SE(3) matrix."""
    assert tuple(matrix.shape) == (4, 4)
    assert points.dim() == 2 and points.shape[1] == 3
    points = _chk(points.detach(), "points")
    out = torch.empty_like(points)
    _lib.check(_lib.lib().gsb_transform_points(points.shape[0], _p(points), _host16(matrix), _p(out), _stream()),
               "transform_points")
    return out


def depth_to_points(depth: Tensor, K: Tensor, include_homogeneous: bool = False, c2w: Tensor | None = None) -> Tensor:
    """geometry.py:138-161: depth [H, W], K [3, 3] -> [H*W, 3] (or [H*W, 4]).  ``c2w`` (extension): apply
    ``transform_points`` in the same pass (dataset.py:349-351 does it as a second step)."""
    depth = _chk(depth.detach(), "depth")
    assert depth.dim() == 2, depth.shape
    H, W = depth.shape
    out = torch.empty((H * W, 3), dtype=torch.float32, device=depth.device)
    _lib.check(_lib.lib().gsb_depth_to_points(W, H, _p(depth), _host16(K[:3, :3]), None if c2w is None else _host16(c2w),
                                              _p(out), _stream()), "depth_to_points")
    if include_homogeneous:
        out = torch.nn.functional.pad(out, (0, 1), value=1)
    return out


@torch.no_grad()
def compute_depth_gt(points: Tensor, rgbs: Tensor | None, Ks: Tensor, c2w: Tensor, height: int, width: int,
                     reference_faithful: bool = True) -> Tensor:
    """geometry.py:69-135: render the expected depth ("ED") of the source cloud from ``c2w`` with opacity 1,
    identity quaternions and KNN scales; returns [height, width].  Colours do not influence the depth channel, so
    ``rgbs`` is accepted for signature parity and ignored."""
    from .rendering import rasterization
    points = _chk(points.detach(), "points")
    N = points.shape[0]
    scales = init_gs_scales(points, reference_faithful=reference_faithful)
    quats = torch.tensor([1.0, 0.0, 0.0, 0.0], device=points.device).repeat(N, 1)
    opac = torch.ones(N, device=points.device)
    colors = torch.zeros((N, 4, 3), device=points.device)
    renders, _, _ = rasterization(means=points, quats=quats, scales=scales, opacities=opac, colors=colors, sh_degree=1,
                                  viewmats=torch.linalg.inv(c2w.float()).to(points.device), Ks=Ks.float().to(points.device),
                                  width=width, height=height, far_plane=1e10, near_plane=1e-2, render_mode="ED",
                                  rasterize_mode="classic")
    return renders.squeeze(0).squeeze(-1)


# ---------------------------------------------------------------------------------------------- PCA normalisation
@torch.no_grad()
def align_principle_axes(point_cloud: Tensor) -> Tensor:
    """src/data/normalize.py:7-50: SE(3) matrix that moves the cloud's per-axis median to the origin and its principal
    axes (eigenvectors of the covariance, largest eigenvalue first, right-handed) onto x, y, z.  The two reductions
    over the N points (median, covariance) run on the GPU (``gsb_points_median_cov``); the 3x3 ``eigh``, the
    determinant flip and the assembly are host code, as cheap as they are in the reference."""
    pts = _chk(point_cloud.detach(), "point_cloud")
    assert pts.dim() == 2 and pts.shape[1] == 3, pts.shape
    med = (C.c_float * 3)()
    cov = (C.c_float * 9)()
    _lib.check(_lib.lib().gsb_points_median_cov(pts.shape[0], _p(pts), med, cov, _stream()), "points_median_cov")
    centroid = torch.tensor(list(med), dtype=torch.float32)
    covariance = torch.tensor(list(cov), dtype=torch.float32).reshape(3, 3)
    eigenvalues, eigenvectors = torch.linalg.eigh(covariance)
    eigenvectors = eigenvectors[:, eigenvalues.argsort(descending=True)]
    if torch.det(eigenvectors) < 0:
        eigenvectors[:, 0] *= -1
    rotation = eigenvectors.t()
    transform = torch.eye(4)
    transform[:3, :3] = rotation
    transform[:3, 3] = -torch.mv(rotation, centroid)
    return transform.to(pts.device)


@torch.no_grad()
def transform_cameras(matrix: Tensor, c2w: Tensor):
    """src/data/normalize.py:75-101: ``matrix @ c2w[n]`` with the rotation block re-normalised by the norm of its first
    row; returns (transformed [N,4,4], scaling [N,1]).  A handful of 4x4 products: plain torch."""
    assert tuple(matrix.shape) == (4, 4)
    assert c2w.dim() == 3 and tuple(c2w.shape[1:]) == (4, 4)
    transformed = torch.einsum("ki,nij->nkj", matrix, c2w)
    scaling = torch.norm(transformed[:, 0, :3], p=2, dim=1, keepdim=True)
    transformed[:, :3, :3] /= scaling.unsqueeze(-1)
    return transformed, scaling


@torch.no_grad()
def normalize_2C(tar_points: Tensor, tar_pose: Tensor, src_points: Tensor, src_pose: Tensor):
    """src/data/normalize.py:104-124 on plain tensors instead of ``RGBDImage`` objects: the transform is computed from the
    TARGET cloud and applied to both clouds and both poses.  Returns (tar_points, tar_pose, src_points, src_pose,
    scale_factor) with scale_factor the target's, as the reference returns it."""
    T = align_principle_axes(tar_points)
    out_t = transform_points(T, tar_points)
    pose_t, scale = transform_cameras(T.to(tar_pose.device), tar_pose.unsqueeze(0))
    out_s = transform_points(T, src_points)
    pose_s, _ = transform_cameras(T.to(src_pose.device), src_pose.unsqueeze(0))
    return out_t, pose_t.squeeze(0), out_s, pose_s.squeeze(0), scale
