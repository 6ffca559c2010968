"""CPU ORACLE — TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED.

This file is the checker, never the product: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  ``gsplatloc_b200`` never does.

It restates, in CPU PyTorch (+ numpy for the integer work), the arithmetic of the path GsplatLoc drives:
``gsplat.rasterization`` as called from ``src/my_gsplat/model.py:195-213`` and
``src/my_gsplat/geometry.py:117-132``, the losses of ``src/my_gsplat/loss.py:10-101`` and the pose
parameterisation / optimiser of ``src/my_gsplat/model.py:27-116`` + ``gs_trainer_total.py:65-267``.

PARITY UNPINNED: the arithmetic lives in an un-vendored submodule, ``thirdparty/gsplat`` =
``SupaVision/gsplat @ 507f26e`` (gsplat 1.3.0, ``.SUBMODULES.json:23-28``), whose directory is empty in
``/root/reference``; ``kornia 0.7.2`` (``poetry.lock:1236-1237``) and ``small_gicp 0.1.1`` are not installed
either and there is no network.  The reference holds no golden vector, fixture or known-answer test for this
path.  So this oracle restates the *published* algorithm of gsplat 1.3.0 / kornia 0.7.2
 and is anchored on the reference's own call sites.  What is checked instead
(tests/test_oracle.py): fp64 autograd against central finite differences for d loss / d(quat, t); forward
invariants (sum(alpha*T) <= 1, offsets monotone, keys sorted, stable ties); convergence of the pose loop to the
known synthetic delta pose.

dtype: every function computes in the dtype of its inputs.  fp64 = gradient truth, fp32 = timed CPU baseline.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn.functional as F

ALPHA_MIN = 1.0 / 255.0      # gsplat rasterize_to_pixels_fwd: skip if alpha < 1/255      (Appendix A.4)
ALPHA_MAX = 0.999            # alpha = min(0.999, opac * exp(-sigma))
T_STOP = 1e-4                # stop the pixel when T * (1 - alpha) <= 1e-4 (entry not composited)
SH_C0 = 0.28209479177387814  # src/my_gsplat/utils.py:53-55
SH_C1 = 0.4886025119029199


# ----------------------------------------------------------------------------------------------------------
# A.2 projection  (gsplat fully_fused_projection_fwd; docs/main.md:155-172 for Sigma' = J W Sigma W^T J^T)
# ----------------------------------------------------------------------------------------------------------
def quat_to_rotmat(q: torch.Tensor) -> torch.Tensor:
    """wxyz quaternion, normalised inside (gsplat quat_to_rotmat; same matrix as kornia's, A.6)."""
    q = q / q.norm(dim=-1, keepdim=True)
    w, x, y, z = q.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(q.shape[:-1] + (3, 3))


# two constants of the absent gsplat fork that neighbouring versions set differently.  The defaults
# are upstream 1.3.0's (what the kernels are built with); the alternatives exist so that the exposure is a NUMBER
# (tests/test_oracle.py::test_exposure_to_the_uncertain_constants) -- the kernels have the same two switches at compile time
# (csrc/common.cuh: -DGSB_RADIUS_FLOOR=0.1f, -DGSB_CLAMP_ASYMMETRIC=1; the environment variables below select the oracle's
# side for a GPU run against such a build).
VARIANT = {"radius_floor": float(os.environ.get("GSB_ORACLE_RADIUS_FLOOR", "0.01")),      # B2: Inria's rasteriser uses 0.1
           "clamp": os.environ.get("GSB_ORACLE_CLAMP", "symmetric")}                      # B1: "asymmetric" from gsplat 1.4.0


def project(means, quats, scales, viewmats, Ks, width, height, eps2d=0.3, near_plane=0.01, far_plane=1e10,
            radius_clip=0.0):
    """-> radii[C,N] int32, means2d[C,N,2], depths[C,N], conics[C,N,3]; culled entries are zero."""
    C, N = viewmats.shape[0], means.shape[0]
    R, t = viewmats[:, :3, :3], viewmats[:, :3, 3]
    pc = torch.einsum("cij,nj->cni", R, means) + t[:, None, :]
    Rq = quat_to_rotmat(quats)
    M = Rq * scales[:, None, :]
    cov = M @ M.transpose(-1, -2)
    covc = torch.einsum("cij,njk,clk->cnil", R, cov, R)
    fx, fy, cx, cy = Ks[:, 0, 0, None], Ks[:, 1, 1, None], Ks[:, 0, 2, None], Ks[:, 1, 2, None]
    x, y, z = pc.unbind(-1)
    tanx, tany = 0.5 * width / fx, 0.5 * height / fy
    if VARIANT["clamp"] == "asymmetric":   # gsplat >= 1.4.0
        lxp, lxn = (width - cx) / fx + 0.3 * tanx, cx / fx + 0.3 * tanx
        lyp, lyn = (height - cy) / fy + 0.3 * tany, cy / fy + 0.3 * tany
    else:
        lxp = lxn = 1.3 * tanx
        lyp = lyn = 1.3 * tany
    rz = 1.0 / z
    tx = z * torch.minimum(lxp, torch.maximum(-lxn, x * rz))
    ty = z * torch.minimum(lyp, torch.maximum(-lyn, y * rz))
    O = torch.zeros_like(z)
    J = torch.stack([fx * rz, O, -fx * tx * rz * rz, O, fy * rz, -fy * ty * rz * rz], -1).reshape(C, N, 2, 3)
    cov2 = J @ covc @ J.transpose(-1, -2)
    m2 = torch.stack([fx * x * rz + cx, fy * y * rz + cy], -1)
    a, b, c = cov2[..., 0, 0] + eps2d, cov2[..., 0, 1], cov2[..., 1, 1] + eps2d
    det = a * c - b * b
    conics = torch.stack([c / det, -b / det, a / det], -1)
    bb = 0.5 * (a + c)
    v1 = bb + torch.sqrt(torch.clamp(bb * bb - det, min=VARIANT["radius_floor"]))
    radius = torch.ceil(3.0 * torch.sqrt(v1))
    valid = (z >= near_plane) & (z <= far_plane) & (det > 0) & (radius > radius_clip)
    valid &= (m2[..., 0] + radius > 0) & (m2[..., 0] - radius < width)
    valid &= (m2[..., 1] + radius > 0) & (m2[..., 1] - radius < height)
    radii = torch.where(valid, radius, torch.zeros_like(radius)).to(torch.int32)
    vm = valid[..., None]
    zero = torch.zeros((), dtype=means.dtype)
    return radii, torch.where(vm, m2, zero), torch.where(valid, z, zero), torch.where(vm, conics, zero)


# ----------------------------------------------------------------------------------------------------------
# A.3 tile intersection, keys, sort, offsets (integer work: numpy, bit-exact contract)
# ----------------------------------------------------------------------------------------------------------
def tile_bbox(means2d: np.ndarray, radii: np.ndarray, tile_size: int, tile_w: int, tile_h: int):
    """fp32 arithmetic exactly as the kernel: (u32)floor/ceil of a negative float saturates to 0."""
    ts = np.float32(tile_size)
    tr = radii.astype(np.float32) / ts
    tx = means2d[..., 0].astype(np.float32) / ts
    ty = means2d[..., 1].astype(np.float32) / ts
    def u(v, hi):
        return np.clip(v, 0, hi).astype(np.int64)
    return (u(np.floor(tx - tr), tile_w), u(np.floor(ty - tr), tile_h),
            u(np.ceil(tx + tr), tile_w), u(np.ceil(ty + tr), tile_h))


def isect_tiles(means2d, radii, depths, tile_size, tile_w, tile_h, sort=True):
    """-> tiles_per_gauss[C,N] int32, isect_ids[M] int64, flatten_ids[M] int32."""
    means2d = np.ascontiguousarray(np.asarray(means2d, dtype=np.float32))
    radii = np.asarray(radii)
    depths = np.ascontiguousarray(np.asarray(depths, dtype=np.float32))
    C, N = radii.shape
    x0, y0, x1, y1 = tile_bbox(means2d, radii, tile_size, tile_w, tile_h)
    tpg = np.where(radii > 0, (y1 - y0) * (x1 - x0), 0).astype(np.int64)
    tile_bits = int(math.floor(math.log2(tile_w * tile_h))) + 1
    cam_bits = int(math.floor(math.log2(C))) + 1
    assert tile_bits + cam_bits <= 32
    flat = tpg.reshape(-1)
    M = int(flat.sum())
    idx = np.repeat(np.arange(C * N, dtype=np.int64), flat)            # emission order: by idx ...
    start = np.repeat(np.cumsum(flat) - flat, flat)
    k = np.arange(M, dtype=np.int64) - start                           # ... then i (rows), then j
    bw = (x1 - x0).reshape(-1)[idx]
    i = y0.reshape(-1)[idx] + k // np.maximum(bw, 1)
    j = x0.reshape(-1)[idx] + k % np.maximum(bw, 1)
    dbits = depths.reshape(-1).view(np.int32).astype(np.int64)[idx]
    ids = ((idx // N) << (32 + tile_bits)) | ((i * tile_w + j) << 32) | dbits
    fl = idx.astype(np.int32)
    if sort:
        nb = 32 + tile_bits + cam_bits
        order = np.argsort(ids & ((1 << nb) - 1), kind="stable")       # stable LSD radix on the low nb bits
        ids, fl = ids[order], fl[order]
    return tpg.astype(np.int32), ids, fl


def isect_offset_encode(isect_ids: np.ndarray, C: int, tile_w: int, tile_h: int) -> np.ndarray:
    """offsets[c,i,j] = first sorted position whose (cam,tile) >= that tile; end of last = M."""
    n_tiles = tile_w * tile_h
    tile_bits = int(math.floor(math.log2(n_tiles))) + 1
    hi = np.asarray(isect_ids, dtype=np.int64) >> 32
    lin = (hi >> tile_bits) * n_tiles + (hi & ((1 << tile_bits) - 1))
    return np.searchsorted(lin, np.arange(C * n_tiles), side="left").astype(np.int32).reshape(C, tile_h, tile_w)


# ----------------------------------------------------------------------------------------------------------
# A.4 / A.5 rasterisation (differentiable through torch autograd; per-tile vectorised over 256 pixels)
# ----------------------------------------------------------------------------------------------------------
def rasterize(means2d, conics, colors, opacities, width, height, tile_size, offsets, flatten_ids,
              tile_chunk: int = 48, record: list | None = None, frozen: list | None = None):
    """Front-to-back compositing with gsplat's skip / stop rules.

    means2d[C,N,2] conics[C,N,3] colors[C,N,D] opacities[C,N]; offsets[C,th,tw] int, flatten_ids[M] int.
    -> render_colors[C,H,W,D], render_alphas[C,H,W,1], last_ids[C,H,W] (int64 numpy; position in the sorted list)

    ``record`` / ``frozen``: save / replay the discrete skip-stop decisions (the compositing function is only
    piecewise smooth; finite-difference checks of its gradient freeze the decisions of the centre point).
    """
    C, N, D = colors.shape
    dt = means2d.dtype
    offsets = np.asarray(offsets).astype(np.int64)
    flatten_ids = np.asarray(flatten_ids).astype(np.int64)
    th, tw = offsets.shape[1:]
    M = flatten_ids.shape[0]
    starts = offsets.reshape(-1)
    ends = np.concatenate([starts[1:], [M]])
    counts = ends - starts
    m2f, cof, colf, opf = means2d.reshape(-1, 2), conics.reshape(-1, 3), colors.reshape(-1, D), opacities.reshape(-1)
    ts = tile_size
    Hp, Wp = th * ts, tw * ts
    out_c = [[None] * (th * tw) for _ in range(C)]
    out_a = [[None] * (th * tw) for _ in range(C)]
    last = np.zeros((C, th * tw, ts * ts), dtype=np.int64)
    ly, lx = torch.meshgrid(torch.arange(ts), torch.arange(ts), indexing="ij")
    ly, lx = ly.reshape(-1).to(dt), lx.reshape(-1).to(dt)
    zc, za = torch.zeros(ts * ts, D, dtype=dt), torch.zeros(ts * ts, dtype=dt)
    order = np.argsort(counts, kind="stable")
    order = order[counts[order] > 0]
    for lo in range(0, len(order), tile_chunk):
        tiles = order[lo:lo + tile_chunk]
        G = int(counts[tiles].max())
        B = len(tiles)
        pos = starts[tiles][:, None] + np.arange(G)[None, :]                    # [B,G] positions in the list
        live = torch.from_numpy(np.arange(G)[None, :] < counts[tiles][:, None])  # [B,G]
        gid = torch.from_numpy(flatten_ids[np.minimum(pos, M - 1)])
        tid = tiles % (th * tw)
        px = (torch.from_numpy((tid % tw) * ts).to(dt)[:, None] + lx[None, :] + 0.5)   # [B,256]
        py = (torch.from_numpy((tid // tw) * ts).to(dt)[:, None] + ly[None, :] + 0.5)
        xy, co = m2f[gid], cof[gid]                                               # [B,G,2], [B,G,3]
        dx = xy[..., 0, None] - px[:, None, :]                                    # [B,G,256]
        dy = xy[..., 1, None] - py[:, None, :]
        sigma = 0.5 * (co[..., 0, None] * dx * dx + co[..., 2, None] * dy * dy) + co[..., 1, None] * dx * dy
        alpha = torch.clamp(opf[gid][..., None] * torch.exp(-sigma), max=ALPHA_MAX)
        valid = (sigma >= 0) & (alpha >= ALPHA_MIN) & live[..., None]
        one_m = torch.where(valid, 1 - alpha, torch.ones((), dtype=dt))
        Tn = torch.cumprod(one_m, dim=1)                                          # T after entry k
        stopped = (valid & (Tn.detach() <= T_STOP)).cumsum(1) > 0                 # stop entry and all after
        inc = valid & ~stopped
        if frozen is not None:
            inc = frozen[lo // tile_chunk]
            one_m = torch.where(inc, 1 - alpha, torch.ones((), dtype=dt))
            Tn = torch.cumprod(one_m, dim=1)
        if record is not None:
            record.append(inc)
        Tp = torch.cat([torch.ones_like(Tn[:, :1]), Tn[:, :-1]], 1)               # T before entry k
        w = torch.where(inc, alpha * Tp, torch.zeros((), dtype=dt))               # vis = alpha * T
        col = torch.einsum("bgp,bgd->bpd", w, colf[gid])
        Tfin = torch.where(inc, one_m, torch.ones((), dtype=dt)).prod(1)          # prod over composited
        alp = 1 - Tfin
        li = torch.where(inc, torch.from_numpy(pos)[..., None], torch.zeros((), dtype=torch.int64)).max(1).values
        for b, t in enumerate(tiles):
            c, tt = divmod(int(t), th * tw)
            out_c[c][tt], out_a[c][tt] = col[b], alp[b]
            last[c, tt] = li[b].numpy()
    def assemble(grid, zero, tail):
        rows = []
        for c in range(C):
            t = torch.stack([g if g is not None else zero for g in grid[c]])      # [T,256,...]
            t = t.reshape((th, tw, ts, ts) + tail).permute((0, 2, 1, 3) + tuple(range(4, 4 + len(tail))))
            rows.append(t.reshape((Hp, Wp) + tail)[:height, :width])
        return torch.stack(rows)
    rc = assemble(out_c, zc, (D,))
    ra = assemble(out_a, za, ())[..., None]
    li = last.reshape(C, th, tw, ts, ts).transpose(0, 1, 3, 2, 4).reshape(C, Hp, Wp)[:, :height, :width]
    return rc, ra, li


# ----------------------------------------------------------------------------------------------------------
# spherical harmonics (degree <= 1 is all GsplatLoc uses: sh_degree=1, model.py:128,201)
# ----------------------------------------------------------------------------------------------------------
def spherical_harmonics(degree: int, dirs, coeffs, masks=None):
    """dirs[...,3] (not normalised), coeffs[...,K,3] -> colours[...,3]  (gsplat compute_sh_fwd, deg<=1)."""
    assert degree <= 1, "oracle restates the degrees the reference path uses"
    res = SH_C0 * coeffs[..., 0, :]
    if degree >= 1:
        d = F.normalize(dirs, dim=-1)
        x, y, z = d[..., 0:1], d[..., 1:2], d[..., 2:3]
        res = res + SH_C1 * (-y * coeffs[..., 1, :] + z * coeffs[..., 2, :] - x * coeffs[..., 3, :])
    if masks is not None:
        res = torch.where(masks[..., None], res, torch.zeros((), dtype=res.dtype))
    return res


# ----------------------------------------------------------------------------------------------------------
# A.1 rasterization(): the boundary the reference calls (model.py:195-213, geometry.py:117-132)
# ----------------------------------------------------------------------------------------------------------
def rasterization(means, quats, scales, opacities, colors, viewmats, Ks, width, height, near_plane=0.01,
                  far_plane=1e10, radius_clip=0.0, eps2d=0.3, sh_degree=None, packed=False, tile_size=16,
                  backgrounds=None, render_mode="RGB", sparse_grad=False, absgrad=False,
                  rasterize_mode="classic", channel_chunk=32, frozen_meta=None):
    assert render_mode in ("RGB", "D", "ED", "RGB+D", "RGB+ED") and backgrounds is None
    assert rasterize_mode == "classic" and not packed
    C, N = viewmats.shape[0], means.shape[0]
    radii, means2d, depths, conics = project(means, quats, scales, viewmats, Ks, width, height, eps2d,
                                             near_plane, far_plane, radius_clip)
    opac = opacities[None].expand(C, N)
    if render_mode in ("D", "ED"):
        cols = depths[..., None]
    else:
        if sh_degree is None:
            cols = colors[None].expand(C, N, -1) if colors.dim() == 2 else colors
        else:
            campos = torch.linalg.inv(viewmats)[:, :3, 3]
            dirs = means[None] - campos[:, None, :]
            cols = torch.clamp_min(spherical_harmonics(sh_degree, dirs, colors[None].expand(C, -1, -1, -1),
                                                       radii > 0) + 0.5, 0.0)
        if render_mode in ("RGB+D", "RGB+ED"):
            cols = torch.cat([cols, depths[..., None]], -1)
    tw, th = math.ceil(width / tile_size), math.ceil(height / tile_size)
    if frozen_meta is None:
        tpg, ids, fl = isect_tiles(means2d.detach().float().numpy(), radii.numpy(),
                                   depths.detach().float().numpy(), tile_size, tw, th)
        offs = isect_offset_encode(ids, C, tw, th)
        rec, frz = [], None
    else:
        tpg, ids, fl, offs = (frozen_meta[k] for k in ("tiles_per_gauss", "isect_ids", "flatten_ids", "isect_offsets"))
        rec, frz = None, frozen_meta["decisions"]
    rc, ra, last = rasterize(means2d, conics, cols, opac, width, height, tile_size, offs, fl, record=rec, frozen=frz)
    if render_mode in ("ED", "RGB+ED"):
        rc = torch.cat([rc[..., :-1], rc[..., -1:] / ra.clamp(min=1e-10)], -1)
    meta = dict(radii=radii, means2d=means2d, depths=depths, conics=conics, opacities=opac, tile_width=tw,
                tile_height=th, tiles_per_gauss=tpg, isect_ids=ids, flatten_ids=fl, isect_offsets=offs,
                width=width, height=height, tile_size=tile_size, n_cameras=C, last_ids=last,
                decisions=rec if rec is not None else frz)
    return rc, ra, meta


# ----------------------------------------------------------------------------------------------------------
# A.7 losses (loss.py:10-101 + kornia 0.7.2 sobel / depth_to_3d_v2) and the trainer's mask
# ----------------------------------------------------------------------------------------------------------
def sobel(img: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """kornia.filters.sobel(normalized=True): img[B,1,H,W] -> sqrt(gx^2 + gy^2 + eps), replicate pad."""
    kx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]], dtype=img.dtype) / 8.0
    k = torch.stack([kx, kx.t()])[:, None]
    g = F.conv2d(F.pad(img, (1, 1, 1, 1), mode="replicate"), k)
    return torch.sqrt(g[:, 0:1] ** 2 + g[:, 1:2] ** 2 + eps)


def depth_to_normal(depth: torch.Tensor, K: torch.Tensor) -> torch.Tensor:
    """geometry.py:164-197: depth[H,W] -> normals[H,W,3] (central differences on replicate-padded points)."""
    H, W = depth.shape
    v, u = torch.meshgrid(torch.arange(H, dtype=depth.dtype), torch.arange(W, dtype=depth.dtype), indexing="ij")
    P = torch.stack([(u - K[0, 2]) / K[0, 0] * depth, (v - K[1, 2]) / K[1, 1] * depth, depth], -1)
    Pp = F.pad(P[None].permute(0, 3, 1, 2), (1, 1, 1, 1), mode="replicate").permute(0, 2, 3, 1)
    dx = Pp[:, 1:-1, 2:, :] - Pp[:, 1:-1, :-2, :]
    dy = Pp[:, 2:, 1:-1, :] - Pp[:, :-2, 1:-1, :]
    return F.normalize(torch.cross(dx, dy, dim=-1), p=2, dim=-1)[0]


def losses(depth, gt, K=None, depth_lambda=0.8, normal_lambda=0.0, normal_dim=1):
    """gs_trainer_total.py:105-150.  depth, gt: [H,W].  normal_dim=1 is the reference's reduction over W
    (loss.py:93); normal_dim=-1 is the per-pixel variant.  -> total, L_d, L_s, L_n"""
    m = (depth != 0).to(depth.dtype)
    A, B = depth * m, gt * m
    Ld = (A - B).abs().mean()
    Ls = (sobel(A[None, None]) - sobel(B[None, None])).abs().mean()
    Ln = torch.zeros((), dtype=depth.dtype)
    if normal_lambda != 0.0:
        Ln = 1 - F.cosine_similarity(depth_to_normal(B, K), depth_to_normal(A, K), dim=normal_dim).mean()
    return depth_lambda * Ld + normal_lambda * Ln + (1 - depth_lambda - normal_lambda) * Ls, Ld, Ls, Ln


# ----------------------------------------------------------------------------------------------------------
# A.6 pose parameterisation (model.py:27-116; kornia quaternion wxyz) and error metrics (eval/utils.py:122-168)
# ----------------------------------------------------------------------------------------------------------
def pose_to_c2w(quat: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    q = quat / quat.norm().clamp_min(1e-12)
    c2w = torch.eye(4, dtype=quat.dtype)
    c2w = torch.cat([torch.cat([quat_to_rotmat(q), t[:, None]], 1), c2w[3:]], 0)
    return c2w


def rotmat_to_quat(R: np.ndarray) -> np.ndarray:
    """kornia rotation_matrix_to_quaternion (branchy trace method), wxyz."""
    R = np.asarray(R, dtype=np.float64)
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = math.sqrt(tr + 1.0 + 1e-8) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2] + 1e-8) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2] + 1e-8) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1] + 1e-8) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    return np.asarray(q)


def pose_errors(c2w_est, c2w_gt):
    """eT = ||dt||, eR = acos((tr(R_est R_gt^T) - 1) / 2) in degrees."""
    c2w_est, c2w_gt = np.asarray(c2w_est, dtype=np.float64), np.asarray(c2w_gt, dtype=np.float64)
    eT = float(np.linalg.norm(c2w_est[:3, 3] - c2w_gt[:3, 3]))
    ct = (np.trace(c2w_est[:3, :3] @ c2w_gt[:3, :3].T) - 1) / 2
    return eT, float(np.degrees(np.arccos(np.clip(ct, -1, 1))))


def pose_loss(quat, t, means, quats, scales, opac, gt, K, W, H, depth_lambda=0.8, normal_lambda=0.0,
              normal_dim=1, frozen_meta=None):
    """One forward of the hot loop (gs_trainer_total.py:95-150) for a single camera; returns
    (total, L_d, L_s, L_n, depth[H,W], alpha[H,W], meta)."""
    view = torch.linalg.inv(pose_to_c2w(quat, t))[None]
    rc, ra, meta = rasterization(means, quats, scales, opac, None, view, K[None], W, H, render_mode="ED",
                                 frozen_meta=frozen_meta)
    d = rc[0, :, :, 0]
    tot, Ld, Ls, Ln = losses(d, gt, K, depth_lambda, normal_lambda, normal_dim)
    return tot, Ld, Ls, Ln, d, ra[0, :, :, 0], meta


def pose_opt_loop(quat0, t0, means, quats, scales, opac, gt, K, W, H, c2w_gt=None, max_steps=1000, n_steps=None,
                  early_stop=True, patience=200, warm=100, lr_q=5e-4, lr_t=1e-3, wd=1e-3, depth_lambda=0.8,
                  normal_lambda=0.0, normal_dim=1, log=None):
    """gs_trainer_total.py:65-267: two Adam optimisers (coupled weight decay), ExponentialLR stepped every
    iteration with gamma = 0.2**(1/max_steps), early stop on the total loss after ``warm`` steps.
    Runs ``n_steps`` (default max_steps) iterations at most.  -> dict(quat, t, steps, best_*, history)"""
    q = quat0.clone().requires_grad_(True)
    t = t0.clone().requires_grad_(True)
    opts = [torch.optim.Adam([q], lr=lr_q, weight_decay=wd), torch.optim.Adam([t], lr=lr_t, weight_decay=wd)]
    gamma = 0.2 ** (1.0 / max_steps)
    best = dict(loss=float("inf"), eT=float("inf"), eR=float("inf"), step=-1)
    counter, hist, stopped = 0, [], False
    steps_done = 0
    for step in range(n_steps or max_steps):
        for o in opts:
            o.zero_grad(set_to_none=True)
        tot, Ld, Ls, Ln, *_ = pose_loss(q, t, means, quats, scales, opac, gt, K, W, H, depth_lambda, normal_lambda,
                                        normal_dim)
        tot.backward()
        eT = eR = float("nan")
        if c2w_gt is not None:
            eT, eR = pose_errors(pose_to_c2w(q.detach(), t.detach()).numpy(), c2w_gt)
        hist.append((float(tot.detach()), float(Ld.detach()), float(Ls.detach()), eT, eR))
        if log:
            log(step, hist[-1], q.grad.clone(), t.grad.clone())
        if early_stop and step > warm:
            if float(tot.detach()) < best["loss"]:
                best.update(loss=float(tot.detach()), eT=eT, eR=eR, step=step)
                counter = 0
            else:
                counter += 1
        if early_stop and counter >= patience:
            stopped = True
            break
        for o in opts:
            o.step()
            for g in o.param_groups:
                g["lr"] *= gamma
        steps_done += 1
    return dict(quat=q.detach(), t=t.detach(), steps=steps_done, stopped=stopped, best=best, history=hist)
