"""Loss functions of ``src/my_gsplat/loss.py`` (same names and argument meaning) on the fused sm_100a loss kernel
(``csrc/loss.cu``): one pass computes the L1 depth term, the Sobel-edge ("silhouette") term and d(loss)/d(depth).

``compute_depth_loss`` (loss.py:10-30) and ``compute_silhouette_loss`` (loss.py:33-59) are autograd functions of
their FIRST argument (the rendered depth); the reference also back-propagates into the observation because its
``TensorWrapper`` marks everything ``requires_grad`` (data/base.py:92-106), a gradient nobody reads — not produced.
"""
from __future__ import annotations

from typing import Tuple

import torch
from torch import Tensor

from . import _lib
from .ops import _chk, _p, _stream


def depth_edge_loss(depth: Tensor, gt: Tensor, depth_lambda: float = 0.8, normal_lambda: float = 0.0,
                    apply_mask: bool = True) -> Tuple[Tensor, Tensor, Tensor]:
    """depth[C,H,W], gt[H,W] | [C,H,W] -> (L_d[C], L_s[C], d(total)/d(depth)[C,H,W]) with
    total = depth_lambda * L_d + (1 - depth_lambda - normal_lambda) * L_s  (gs_trainer_total.py:145-150) and the
    trainer's mask ``depth != 0`` (:109) applied to both images when ``apply_mask``."""
    assert depth.dim() == 3, depth.shape
    Cn, H, W = depth.shape
    per_cam = gt.dim() == 3
    assert tuple(gt.shape[-2:]) == (H, W) and (not per_cam or gt.shape[0] == Cn), gt.shape
    depth, gt = _chk(depth, "depth"), _chk(gt, "gt")
    inv_p = 1.0 / float(H * W)
    sums = torch.empty((Cn, 2), dtype=torch.float32, device=depth.device)
    v = torch.empty_like(depth)
    ws = torch.empty(2 * Cn * ((W + 31) // 32) * ((H + 7) // 8), dtype=torch.float32, device=depth.device)
    _lib.check(_lib.lib().gsb_depth_edge_loss(Cn, W, H, _p(depth), _p(gt), int(per_cam), int(apply_mask),
                                              depth_lambda * inv_p, (1.0 - depth_lambda - normal_lambda) * inv_p,
                                              _p(sums), _p(v), _p(ws), _stream()), "depth_edge_loss")
    return sums[:, 0] * inv_p, sums[:, 1] * inv_p, v


class _FusedL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, A, B, which):
        lead = A.shape
        H, W = (lead[-3], lead[-2]) if (A.dim() == 4 and lead[-1] == 1) else (lead[-2], lead[-1])
        a = A.reshape(-1, H, W)
        b = B.reshape(-1, H, W)
        n = a.shape[0]
        Ld, Ls, v = depth_edge_loss(a, b, depth_lambda=1.0 if which == 0 else 0.0, apply_mask=False)
        ctx.save_for_backward(v)
        ctx.shape, ctx.n = lead, n
        return (Ld if which == 0 else Ls).mean()

    @staticmethod
    def backward(ctx, g):
        (v,) = ctx.saved_tensors
        return (v * (g / ctx.n)).reshape(ctx.shape), None, None


def compute_depth_loss(depth_A: Tensor, depth_B: Tensor, *, loss_type: str = "l1") -> Tensor:
    """``F.l1_loss(depth_A, depth_B)`` (loss.py:10-30); shape (*, H, W, 1) or (*, H, W)."""
    if loss_type != "l1":
        raise ValueError("only the 'l1' loss the shipped trainer uses is built (gs_trainer_total.py:126-130)")
    assert depth_A.shape == depth_B.shape, (depth_A.shape, depth_B.shape)
    return _FusedL1.apply(depth_A, depth_B.detach(), 0)


def compute_silhouette_loss(depth_A: Tensor, depth_B: Tensor, *, loss_type: str = "l1") -> Tensor:
    """``F.l1_loss(sobel(depth_A), sobel(depth_B))`` with kornia's normalised Sobel magnitude (loss.py:33-59);
    shape (B, H, W, 1)."""
    if loss_type != "l1":
        raise ValueError("only the 'l1' loss the shipped trainer uses is built (gs_trainer_total.py:133-137)")
    assert depth_A.dim() == 4 and depth_B.dim() == 4
    return _FusedL1.apply(depth_A, depth_B.detach(), 1)
