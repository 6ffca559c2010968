"""``rasterization()`` with gsplat 1.3.0's call signature — the function GsplatLoc imports
(``from gsplat import rasterization``, ``src/my_gsplat/model.py:5,195-213`` and ``geometry.py:4,117-132``).

Same keywords, same return triple ``(render_colors[C,H,W,D], render_alphas[C,H,W,1], meta)``, differentiable with
respect to ``viewmats`` (mandatory for the pose path) and, lazily, ``means / quats / scales / opacities / colors``.
Every stage is a hand-written sm_100a kernel reached through the C ABI (``ops.py``).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

from . import ops


def rasterization(means: Tensor, quats: Tensor, scales: Tensor, opacities: Tensor, colors: Tensor, viewmats: Tensor,
                  Ks: Tensor, width: int, height: int, near_plane: float = 0.01, far_plane: float = 1e10,
                  radius_clip: float = 0.0, eps2d: float = 0.3, sh_degree: Optional[int] = None, packed: bool = True,
                  tile_size: int = 16, backgrounds: Optional[Tensor] = None, render_mode: str = "RGB",
                  sparse_grad: bool = False, absgrad: bool = False, rasterize_mode: str = "classic",
                  channel_chunk: int = 32, distributed: bool = False, ortho: bool = False,
                  covars: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Dict]:
    N = means.shape[0]
    C = viewmats.shape[0]
    assert means.shape == (N, 3), means.shape
    assert quats.shape == (N, 4), quats.shape
    assert scales.shape == (N, 3), scales.shape
    assert opacities.shape == (N,), opacities.shape
    assert viewmats.shape == (C, 4, 4), viewmats.shape
    assert Ks.shape == (C, 3, 3), Ks.shape
    assert render_mode in ["RGB", "D", "ED", "RGB+D", "RGB+ED"], render_mode
    assert rasterize_mode == "classic", "GsplatLoc renders with rasterize_mode='classic' (model.py:212)"
    assert covars is None and not distributed and not ortho, "not on GsplatLoc's path"
    assert not absgrad and not sparse_grad, "GsplatLoc passes absgrad=False, sparse_grad=False (model.py:123-125)"
    # ``packed`` only changes the memory layout of intermediates in gsplat; results are identical. The B200 path
    # always uses its own dense layout, so the flag is accepted and ignored (GsplatLoc passes packed=False).
    need_rgb = render_mode not in ("D", "ED")
    if need_rgb:
        if sh_degree is None:
            assert (colors.dim() == 2 and colors.shape[0] == N) or (colors.dim() == 3 and colors.shape[:2] == (C, N)), colors.shape
        else:
            assert (colors.dim() == 3 and colors.shape[0] == N and colors.shape[2] == 3) or (
                colors.dim() == 4 and colors.shape[:2] == (C, N) and colors.shape[3] == 3), colors.shape
            assert (sh_degree + 1) ** 2 <= colors.shape[-2], colors.shape

    radii, means2d, depths, conics, _ = ops.fully_fused_projection(
        means, None, quats, scales, viewmats, Ks, width, height, eps2d=eps2d, near_plane=near_plane,
        far_plane=far_plane, radius_clip=radius_clip)
    opac = opacities.repeat(C, 1)  # [C, N]
    meta = {"camera_ids": None, "gaussian_ids": None, "radii": radii, "means2d": means2d, "depths": depths,
            "conics": conics, "opacities": opac}

    if need_rgb:
        if sh_degree is None:
            cols = colors.expand(C, -1, -1) if colors.dim() == 2 else colors
        else:
            camtoworlds = torch.inverse(viewmats)
            dirs = means[None, :, :] - camtoworlds[:, None, :3, 3]
            shs = colors.expand(C, -1, -1, -1) if colors.dim() == 3 else colors
            cols = ops.spherical_harmonics(sh_degree, dirs, shs, masks=radii > 0)
            cols = torch.clamp_min(cols + 0.5, 0.0)
        if render_mode in ("RGB+D", "RGB+ED"):
            cols = torch.cat((cols, depths[..., None]), dim=-1)
            if backgrounds is not None:
                backgrounds = torch.cat([backgrounds, backgrounds.new_zeros(C, 1)], dim=-1)
    else:
        cols = depths[..., None]
        if backgrounds is not None:
            backgrounds = backgrounds.new_zeros(C, 1)

    tile_width, tile_height = ops.tile_grid(width, height, tile_size)
    tiles_per_gauss, isect_ids, flatten_ids = ops.isect_tiles(means2d, radii, depths, tile_size, tile_width,
                                                              tile_height, n_cameras=C)
    isect_offsets = ops.isect_offset_encode(isect_ids, C, tile_width, tile_height)
    meta.update({"tile_width": tile_width, "tile_height": tile_height, "tiles_per_gauss": tiles_per_gauss,
                 "isect_ids": isect_ids, "flatten_ids": flatten_ids, "isect_offsets": isect_offsets, "width": width,
                 "height": height, "tile_size": tile_size, "n_cameras": C})

    D = cols.shape[-1]
    if D > channel_chunk or D > 8:
        step = min(channel_chunk, 8)
        parts, ra = [], None
        for lo in range(0, D, step):
            bg = backgrounds[..., lo:lo + step] if backgrounds is not None else None
            rc_i, ra = ops.rasterize_to_pixels(means2d, conics, cols[..., lo:lo + step].contiguous(), opac, width,
                                               height, tile_size, isect_offsets, flatten_ids, backgrounds=bg)
            parts.append(rc_i)
        render_colors, render_alphas = torch.cat(parts, dim=-1), ra
    else:
        render_colors, render_alphas = ops.rasterize_to_pixels(means2d, conics, cols.contiguous(), opac, width,
                                                               height, tile_size, isect_offsets, flatten_ids,
                                                               backgrounds=backgrounds)
    if render_mode in ["ED", "RGB+ED"]:
        render_colors = torch.cat([render_colors[..., :-1],
                                   render_colors[..., -1:] / render_alphas.clamp(min=1e-10)], dim=-1)
    return render_colors, render_alphas, meta
