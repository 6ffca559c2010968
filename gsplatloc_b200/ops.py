"""Op-level boundary: same names, argument meaning and error behaviour as ``gsplat.cuda._wrapper``
(``fully_fused_projection`` IDX:14351, ``isect_tiles`` IDX:14360, ``isect_offset_encode`` IDX:14369,
``rasterize_to_pixels`` IDX:14378, ``spherical_harmonics`` IDX:14306 of
``.vscode/PythonImportHelper-v2-Completion.json`` in the reference), implemented by hand-written sm_100a kernels
behind the C ABI.  Tensors are borrowed (CUDA, fp32, made contiguous), outputs are allocated on the inputs' device,
work is enqueued on the current CUDA stream.  No CPU path: a non-CUDA tensor raises.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib

_RASTER_CHANNELS = (1, 2, 3, 4, 5, 8)


def _chk(t: Optional[Tensor], name: str, dtype=torch.float32) -> Optional[Tensor]:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (gsplatloc_b200 has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _p(t: Optional[Tensor]):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------------------------------------ projection
class _FullyFusedProjection(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, quats, scales, viewmats, Ks, width, height, eps2d, near_plane, far_plane, radius_clip):
        N, Cn = means.shape[0], viewmats.shape[0]
        radii = torch.empty((Cn, N), dtype=torch.int32, device=means.device)
        means2d = torch.empty((Cn, N, 2), dtype=torch.float32, device=means.device)
        depths = torch.empty((Cn, N), dtype=torch.float32, device=means.device)
        conics = torch.empty((Cn, N, 3), dtype=torch.float32, device=means.device)
        _lib.check(_lib.lib().gsb_projection_fwd(Cn, N, _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks), width,
                                                 height, eps2d, near_plane, far_plane, radius_clip, _p(radii),
                                                 _p(means2d), _p(depths), _p(conics), _stream()), "projection_fwd")
        ctx.save_for_backward(means, quats, scales, viewmats, Ks, radii)
        ctx.dims = (width, height, eps2d)
        ctx.mark_non_differentiable(radii)
        return radii, means2d, depths, conics

    @staticmethod
    def backward(ctx, v_radii, v_means2d, v_depths, v_conics):
        means, quats, scales, viewmats, Ks, radii = ctx.saved_tensors
        width, height, eps2d = ctx.dims
        N, Cn = means.shape[0], viewmats.shape[0]
        need = ctx.needs_input_grad
        v_means = torch.zeros_like(means) if need[0] else None
        v_quats = torch.zeros_like(quats) if need[1] else None
        v_scales = torch.zeros_like(scales) if need[2] else None
        v_view = torch.zeros_like(viewmats) if need[3] else None
        if any(need[:4]):
            # contiguous copies are bound to names: a temporary would be freed (and its block possibly reused by the
            # next temporary) before the kernel that reads it is enqueued
            vm2, vdp, vcn = v_means2d.contiguous(), v_depths.contiguous(), v_conics.contiguous()
            _lib.check(_lib.lib().gsb_projection_bwd(
                Cn, N, _p(means), _p(quats), _p(scales), _p(viewmats), _p(Ks), width, height, eps2d, _p(radii),
                _p(vm2), _p(vdp), _p(vcn), _p(v_means),
                _p(v_quats), _p(v_scales), _p(v_view), _stream()), "projection_bwd")
        return v_means, v_quats, v_scales, v_view, None, None, None, None, None, None, None


def fully_fused_projection(means: Tensor, covars: Optional[Tensor], quats: Optional[Tensor],
                           scales: Optional[Tensor], viewmats: Tensor, Ks: Tensor, width: int, height: int,
                           eps2d: float = 0.3, near_plane: float = 0.01, far_plane: float = 1e10,
                           radius_clip: float = 0.0, packed: bool = False, sparse_grad: bool = False,
                           calc_compensations: bool = False):
    """-> (radii[C,N] i32, means2d[C,N,2], depths[C,N], conics[C,N,3], compensations=None)."""
    assert covars is None, "covars input is not built: GsplatLoc passes quats/scales (model.py:195-213)"
    assert not packed and not sparse_grad, "packed/sparse_grad are not used by GsplatLoc (model.py:123-125)"
    assert not calc_compensations, "rasterize_mode='classic' only (model.py:212)"
    N = means.shape[0]
    assert means.shape == (N, 3) and quats.shape == (N, 4) and scales.shape == (N, 3), (means.shape, quats.shape)
    Cn = viewmats.shape[0]
    assert viewmats.shape == (Cn, 4, 4) and Ks.shape == (Cn, 3, 3), (viewmats.shape, Ks.shape)
    means, quats, scales, viewmats, Ks = (_chk(t, n) for t, n in ((means, "means"), (quats, "quats"),
                                                                  (scales, "scales"), (viewmats, "viewmats"),
                                                                  (Ks, "Ks")))
    radii, means2d, depths, conics = _FullyFusedProjection.apply(means, quats, scales, viewmats, Ks, int(width),
                                                                 int(height), float(eps2d), float(near_plane),
                                                                 float(far_plane), float(radius_clip))
    return radii, means2d, depths, conics, None


# ------------------------------------------------------------------------------------------------ intersection
@torch.no_grad()
def isect_tiles(means2d: Tensor, radii: Tensor, depths: Tensor, tile_size: int, tile_width: int, tile_height: int,
                sort: bool = True, packed: bool = False, n_cameras: Optional[int] = None,
                camera_ids: Optional[Tensor] = None, gaussian_ids: Optional[Tensor] = None,
                method: str = "tile") -> Tuple[Tensor, Tensor, Tensor]:
    """-> (tiles_per_gauss[C,N] i32, isect_ids[M] i64, flatten_ids[M] i32).  ``method``: "tile" (per-tile
    shared-memory sort, default) or "radix" (global stable radix sort, the reference's shape); identical outputs."""
    assert not packed, "packed mode is not used by GsplatLoc"
    Cn, N = radii.shape
    assert means2d.shape == (Cn, N, 2) and depths.shape == (Cn, N), (means2d.shape, depths.shape)
    means2d, depths = _chk(means2d, "means2d"), _chk(depths, "depths")
    radii = _chk(radii, "radii", torch.int32)
    dev = means2d.device
    n_bins = Cn * tile_width * tile_height
    tpg = torch.empty((Cn, N), dtype=torch.int32, device=dev)
    offsets = torch.empty(n_bins + 1, dtype=torch.int32, device=dev)
    counts_ws = torch.empty(2 * n_bins + 4, dtype=torch.int32, device=dev)
    m = _lib.c_i64(0)
    L = _lib.lib()
    _lib.check(L.gsb_isect_count(Cn, N, _p(means2d), _p(radii), tile_size, tile_width, tile_height, _p(tpg),
                                 _p(offsets), _p(counts_ws), m, _stream()), "isect_count")
    M = int(m.value)
    meth = {"tile": 0, "radix": 1}[method] if sort else 1
    isect_ids = torch.empty(M, dtype=torch.int64, device=dev)
    flatten_ids = torch.empty(M, dtype=torch.int32, device=dev)
    ws_bytes = L.gsb_isect_emit_workspace(Cn, N, M, tile_width, tile_height, meth)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    _lib.check(L.gsb_isect_emit(Cn, N, _p(means2d), _p(radii), _p(depths), _p(tpg), _p(offsets), tile_size,
                                tile_width, tile_height, int(bool(sort)), meth, M, _p(isect_ids), _p(flatten_ids),
                                _p(ws), ws_bytes, _stream()), "isect_emit")
    return tpg, isect_ids, flatten_ids


@torch.no_grad()
def isect_offset_encode(isect_ids: Tensor, n_cameras: int, tile_width: int, tile_height: int) -> Tensor:
    """sorted isect_ids[M] -> offsets[C, tile_height, tile_width] i32."""
    isect_ids = _chk(isect_ids, "isect_ids", torch.int64)
    offsets = torch.empty((n_cameras, tile_height, tile_width), dtype=torch.int32, device=isect_ids.device)
    _lib.check(_lib.lib().gsb_isect_offset_encode(isect_ids.numel(), _p(isect_ids), n_cameras, tile_width,
                                                  tile_height, _p(offsets), _stream()), "isect_offset_encode")
    return offsets


# ------------------------------------------------------------------------------------------------ rasterisation
class _RasterizeToPixels(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means2d, conics, colors, opacities, backgrounds, width, height, tile_size, isect_offsets,
                flatten_ids):
        Cn, N, D = colors.shape
        dev = means2d.device
        render_colors = torch.empty((Cn, height, width, D), dtype=torch.float32, device=dev)
        render_alphas = torch.empty((Cn, height, width, 1), dtype=torch.float32, device=dev)
        last_ids = torch.empty((Cn, height, width), dtype=torch.int32, device=dev)
        _lib.check(_lib.lib().gsb_rasterize_fwd(Cn, N, D, _p(means2d), _p(conics), _p(colors), _p(opacities),
                                                _p(backgrounds), width, height, tile_size, _p(isect_offsets),
                                                _p(flatten_ids), flatten_ids.numel(), _p(render_colors),
                                                _p(render_alphas), _p(last_ids), _stream()), "rasterize_fwd")
        ctx.save_for_backward(means2d, conics, colors, opacities, backgrounds, isect_offsets, flatten_ids,
                              render_alphas, last_ids)
        ctx.dims = (width, height, tile_size)
        return render_colors, render_alphas

    @staticmethod
    def backward(ctx, v_render_colors, v_render_alphas):
        (means2d, conics, colors, opacities, backgrounds, isect_offsets, flatten_ids, render_alphas,
         last_ids) = ctx.saved_tensors
        width, height, tile_size = ctx.dims
        Cn, N, D = colors.shape
        v_means2d, v_conics = torch.zeros_like(means2d), torch.zeros_like(conics)
        v_colors, v_opac = torch.zeros_like(colors), torch.zeros_like(opacities)
        vrc, vra = v_render_colors.contiguous(), v_render_alphas.contiguous()   # named: alive until the launch
        _lib.check(_lib.lib().gsb_rasterize_bwd(
            Cn, N, D, _p(means2d), _p(conics), _p(colors), _p(opacities), _p(backgrounds), width, height, tile_size,
            _p(isect_offsets), _p(flatten_ids), flatten_ids.numel(), _p(render_alphas), _p(last_ids),
            _p(vrc), _p(vra), _p(v_means2d), _p(v_conics),
            _p(v_colors), _p(v_opac), _stream()), "rasterize_bwd")
        v_bg = None
        if backgrounds is not None and ctx.needs_input_grad[4]:
            v_bg = (v_render_colors * (1.0 - render_alphas)).sum(dim=(1, 2))
        return v_means2d, v_conics, v_colors, v_opac, v_bg, None, None, None, None, None


def rasterize_to_pixels(means2d: Tensor, conics: Tensor, colors: Tensor, opacities: Tensor, image_width: int,
                        image_height: int, tile_size: int, isect_offsets: Tensor, flatten_ids: Tensor,
                        backgrounds: Optional[Tensor] = None, packed: bool = False,
                        absgrad: bool = False) -> Tuple[Tensor, Tensor]:
    """-> (render_colors[C,H,W,D], render_alphas[C,H,W,1])."""
    assert not packed and not absgrad, "packed/absgrad are not used by GsplatLoc (model.py:123-125)"
    Cn, N = means2d.shape[:2]
    assert means2d.shape == (Cn, N, 2) and conics.shape == (Cn, N, 3) and opacities.shape == (Cn, N)
    assert colors.shape[:2] == (Cn, N), colors.shape
    th, tw = isect_offsets.shape[1:]
    assert th * tile_size >= image_height and tw * tile_size >= image_width, "tile grid does not cover the image"
    D = colors.shape[-1]
    Dp = next((d for d in _RASTER_CHANNELS if d >= D), None)
    if Dp is None:
        raise RuntimeError(f"rasterize_to_pixels: {D} channels not supported (max {_RASTER_CHANNELS[-1]})")
    if Dp != D:
        colors = torch.cat([colors, colors.new_zeros(Cn, N, Dp - D)], -1)
        if backgrounds is not None:
            backgrounds = torch.cat([backgrounds, backgrounds.new_zeros(Cn, Dp - D)], -1)
    means2d, conics, colors, opacities = (_chk(t, n) for t, n in ((means2d, "means2d"), (conics, "conics"),
                                                                  (colors, "colors"), (opacities, "opacities")))
    backgrounds = _chk(backgrounds, "backgrounds")
    isect_offsets = _chk(isect_offsets, "isect_offsets", torch.int32)
    flatten_ids = _chk(flatten_ids, "flatten_ids", torch.int32)
    rc, ra = _RasterizeToPixels.apply(means2d, conics, colors, opacities, backgrounds, int(image_width),
                                      int(image_height), int(tile_size), isect_offsets, flatten_ids)
    return (rc[..., :D] if Dp != D else rc), ra


# ------------------------------------------------------------------------------------------------ SH
class _SphericalHarmonics(torch.autograd.Function):
    @staticmethod
    def forward(ctx, degree, dirs, coeffs, masks):
        n, K = dirs.shape[0], coeffs.shape[1]
        colors = torch.empty((n, 3), dtype=torch.float32, device=dirs.device)
        _lib.check(_lib.lib().gsb_sh_fwd(n, K, degree, _p(dirs), _p(coeffs), _p(masks), _p(colors), _stream()), "sh_fwd")
        ctx.save_for_backward(dirs, coeffs, masks)
        ctx.degree = degree
        return colors

    @staticmethod
    def backward(ctx, v_colors):
        dirs, coeffs, masks = ctx.saved_tensors
        n, K = dirs.shape[0], coeffs.shape[1]
        v_coeffs = torch.zeros_like(coeffs)
        v_dirs = torch.zeros_like(dirs) if ctx.needs_input_grad[1] else None
        vcol = v_colors.contiguous()   # named: alive until the launch
        _lib.check(_lib.lib().gsb_sh_bwd(n, K, ctx.degree, _p(dirs), _p(coeffs), _p(masks),
                                         _p(vcol), _p(v_coeffs), _p(v_dirs), _stream()), "sh_bwd")
        return None, v_dirs, v_coeffs, None


def spherical_harmonics(degrees_to_use: int, dirs: Tensor, coeffs: Tensor, masks: Optional[Tensor] = None) -> Tensor:
    """dirs[...,3], coeffs[...,K,3], masks[...] -> colours[...,3]; degrees 0 and 1."""
    assert (degrees_to_use + 1) ** 2 <= coeffs.shape[-2], coeffs.shape
    assert dirs.shape[:-1] == coeffs.shape[:-2] and dirs.shape[-1] == 3 and coeffs.shape[-1] == 3
    if degrees_to_use > 1:
        raise RuntimeError("spherical_harmonics: only degrees 0 and 1 are built (GsplatLoc uses sh_degree=1)")
    lead = dirs.shape[:-1]
    d = _chk(dirs.reshape(-1, 3), "dirs")
    c = _chk(coeffs.reshape(-1, coeffs.shape[-2], 3), "coeffs")
    m = None if masks is None else _chk(masks.reshape(-1).to(torch.uint8), "masks", torch.uint8)
    return _SphericalHarmonics.apply(int(degrees_to_use), d, c, m).reshape(lead + (3,))


def tile_grid(width: int, height: int, tile_size: int = 16) -> Tuple[int, int]:
    return math.ceil(width / float(tile_size)), math.ceil(height / float(tile_size))
