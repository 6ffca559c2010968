"""Shared builders for the parity tests: small seeded scenes in numpy / torch."""
import numpy as np
import torch

from gsplatloc_b200 import synthetic as syn
from oracle import gsplat_oracle as orc

SMALL_CAM = dict(W=160, H=96, fx=80.0, fy=80.0, cx=79.5, cy=47.5)
MID_CAM = dict(W=320, H=192, fx=160.0, fy=160.0, cx=159.5, cy=95.5)


def frame(cam=SMALL_CAM, stride=2, seed=42, **kw):
    return syn.make_frame_pair(cam, stride=stride, seed=seed, **kw)


def map_tensors(fp, dtype=torch.float32, device="cpu"):
    N = fp.tar_points.shape[0]
    means = torch.tensor(fp.tar_points, dtype=dtype, device=device)
    quats = torch.tensor([[1.0, 0, 0, 0]], dtype=dtype, device=device).repeat(N, 1)
    scales = torch.tensor(fp.scales, dtype=dtype, device=device)
    opac = torch.ones(N, dtype=dtype, device=device)
    return means, quats, scales, opac


def offgrid_pose(fp, dtype=torch.float64):
    """Initial pose nudged off the back-projection lattice (at the lattice pose, tile boxes sit exactly on tile
    borders and the discrete tile assignment flips with rounding noise)."""
    q = torch.tensor(orc.rotmat_to_quat(fp.tar_c2w[:3, :3]), dtype=dtype) + torch.tensor([1e-3, 2e-3, -3e-3, 1e-3], dtype=dtype)
    t = torch.tensor(fp.tar_c2w[:3, 3], dtype=dtype) + torch.tensor([3e-3, -2e-3, 4e-3], dtype=dtype)
    return q, t


def oracle_gt_depth(fp, dtype=torch.float32):
    """Splat-rendered observation of the source frame from its own pose (src/data/dataset.py:362-372)."""
    N = fp.src_points.shape[0]
    means = torch.tensor(fp.src_points, dtype=dtype)
    quats = torch.tensor([[1.0, 0, 0, 0]], dtype=dtype).repeat(N, 1)
    view = torch.linalg.inv(torch.tensor(fp.src_c2w, dtype=dtype))[None]
    with torch.no_grad():
        rc, _, _ = orc.rasterization(means, quats, torch.tensor(fp.src_scales, dtype=dtype), torch.ones(N, dtype=dtype),
                                     None, view, torch.tensor(fp.K, dtype=dtype)[None], fp.W, fp.H, render_mode="ED")
    return rc[0, :, :, 0]


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
