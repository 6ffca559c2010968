"""Loss functions of ``src/my_gsplat/loss.py`` (same names and argument meaning) on the fused sm_100a loss kernel
(``csrc/loss.cu``): one pass computes the L1 depth term, the Sobel-edge ("silhouette") term and d(loss)/d(depth).

``compute_depth_loss`` (loss.py:10-30) and ``compute_silhouette_loss`` (loss.py:33-59) are autograd functions of
their FIRST argument (the rendered depth); the reference also back-propagates into the observation because its
``TensorWrapper`` marks everything ``requires_grad`` (data/base.py:92-106), a gradient nobody reads — not produced.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib
from .ops import _chk, _p, _stream


NORMAL_MODES = {"off": 0, "pixel": 1, "reference": 2}


def depth_edge_normal_loss(depth: Tensor, gt: Tensor, K: Optional[Tensor] = None, depth_lambda: float = 0.8,
                           normal_lambda: float = 0.0, normal_mode: str = "reference", apply_mask: bool = True,
                           weights: Optional[Tuple[float, float, float]] = None):
    """depth[C,H,W], gt[H,W] | [C,H,W], K[3,3] | [C,3,3] -> (L_d[C], L_s[C], L_n[C], d(total)/d(depth)[C,H,W]) with
    total = depth_lambda * L_d + normal_lambda * L_n + (1 - depth_lambda - normal_lambda) * L_s
    (gs_trainer_total.py:145-150) and the trainer's mask ``depth != 0`` (:109) applied to both images when
    ``apply_mask``.  ``weights`` = (w_d, w_s, w_n) overrides the three factors."""
    assert depth.dim() == 3, depth.shape
    Cn, H, W = depth.shape
    per_cam = gt.dim() == 3
    assert tuple(gt.shape[-2:]) == (H, W) and (not per_cam or gt.shape[0] == Cn), gt.shape
    depth, gt = _chk(depth, "depth"), _chk(gt, "gt")
    w_d, w_s, w_n = weights if weights is not None else (depth_lambda, 1.0 - depth_lambda - normal_lambda, normal_lambda)
    mode = NORMAL_MODES[normal_mode] if w_n != 0.0 else 0
    Ks = None
    if mode:
        assert K is not None, "the normal term needs the intrinsics K"
        Ks = _chk(K[..., :3, :3].reshape(-1, 3, 3).expand(Cn, 3, 3).contiguous(), "K")
    inv_p = 1.0 / float(H * W)
    L = _lib.lib()
    sums = torch.empty((Cn, 3), dtype=torch.float32, device=depth.device)
    v = torch.empty_like(depth)
    ws = torch.empty(L.gsb_loss_workspace_floats(Cn, W, H), dtype=torch.float32, device=depth.device)
    _lib.check(L.gsb_depth_edge_normal_loss(Cn, W, H, _p(depth), _p(gt), int(per_cam), int(apply_mask), _p(Ks),
                                            w_d * inv_p, w_s * inv_p, w_n, mode, _p(sums), _p(v), _p(ws), _stream()),
               "depth_edge_normal_loss")
    Ln = 1.0 - sums[:, 2] * (inv_p if mode == 1 else 1.0 / (3.0 * H)) if mode else torch.zeros_like(sums[:, 2])
    return sums[:, 0] * inv_p, sums[:, 1] * inv_p, Ln, v


def depth_edge_loss(depth: Tensor, gt: Tensor, depth_lambda: float = 0.8, normal_lambda: float = 0.0,
                    apply_mask: bool = True) -> Tuple[Tensor, Tensor, Tensor]:
    """The two terms the shipped trainer uses (normal_lambda only rescales the edge weight, base.py:27-28)."""
    Ld, Ls, _, v = depth_edge_normal_loss(depth, gt, None, weights=(depth_lambda, 1.0 - depth_lambda - normal_lambda, 0.0),
                                          apply_mask=apply_mask)
    return Ld, Ls, v


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


class _FusedNormal(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth_rendered, depth_real, K, mode):
        _, _, Ln, v = depth_edge_normal_loss(depth_rendered[None], depth_real[None], K, weights=(0.0, 0.0, 1.0),
                                             normal_mode=mode, apply_mask=False)
        ctx.save_for_backward(v)
        return Ln[0]

    @staticmethod
    def backward(ctx, g):
        (v,) = ctx.saved_tensors
        return v[0] * g, None, None, None


def compute_normal_consistency_loss(depth_real: Tensor, depth_rendered: Tensor, *, K: Tensor,
                                    loss_type: str = "cosine", reduce: str = "reference") -> Tensor:
    """``1 - cosine_similarity(normal_real, normal_rendered, dim=1).mean()`` (loss.py:62-101), differentiable in
    ``depth_rendered``.  Shapes [H,W] or [1,H,W]; K is the reference's 4x4 (or a 3x3) intrinsics matrix.
    ``reduce="reference"`` keeps the reference's reduction over dim=1 of the [H,W,3] normal maps (the image width,
    SURVEY.md Appendix C #2); ``reduce="pixel"`` is the per-pixel cosine the paper describes."""
    if loss_type != "cosine":
        raise ValueError("only the 'cosine' loss (the reference's default, loss.py:67) is built")
    if depth_real.dim() == 3:
        depth_real = depth_real.squeeze(0)
    if depth_rendered.dim() == 3:
        depth_rendered = depth_rendered.squeeze(0)
    assert depth_real.shape == depth_rendered.shape and depth_real.dim() == 2
    return _FusedNormal.apply(depth_rendered, depth_real.detach(), K.detach(), reduce)
