"""Seeded synthetic RGB-D frame pairs shaped like the reference's datasets.

The reference reads Replica / TUM depth PNGs from disk (``src/data/dataset.py:78-321``)
and back-projects them (``src/data/Image.py:25-35``).  Neither dataset is available here,
so this module ray-casts a procedural room analytically and produces the same objects the
reference's ``Parser.__getitem__`` hands to the trainer (``src/data/dataset.py:345-383``):

* ``tar_points``  – world-space points of frame *i* (they become the Gaussians),
* ``tar_c2w``     – pose of frame *i* (initial guess of the optimisation),
* ``src_c2w``     – pose of frame *i+1* (ground truth the optimiser should reach),
* ``src_depth``   – depth image of frame *i+1* (the observation),
* ``scales``      – isotropic scale per point: RMS distance to the 4 nearest neighbours
                    (``src/my_gsplat/geometry.py:44-66`` with ``k=5`` incl. self; true RMS
                    distance).

Everything is numpy on the host; nothing here is on the timed path.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

# datasets/Replica/cam_params.json:3-8
REPLICA = dict(W=1200, H=680, fx=600.0, fy=600.0, cx=599.5, cy=339.5)
# TUM fr1 defaults (cam_params.json for TUM is not in the reference tree)
TUM = dict(W=640, H=480, fx=525.0, fy=525.0, cx=319.5, cy=239.5)
HD = dict(W=1920, H=1080, fx=960.0, fy=960.0, cx=959.5, cy=539.5)


@dataclass
class Scene:
    half: np.ndarray          # room half extents (3,)
    sph_c: np.ndarray         # [S,3]
    sph_r: np.ndarray         # [S]
    box_lo: np.ndarray        # [B,3]
    box_hi: np.ndarray        # [B,3]


@dataclass
class FramePair:
    """What the reference calls ``AlignData`` (``src/data/base.py:109-125``), numpy, fp32."""
    tar_points: np.ndarray    # [N,3] world
    scales: np.ndarray        # [N,3]
    tar_c2w: np.ndarray       # [4,4]
    src_c2w: np.ndarray       # [4,4]
    src_depth: np.ndarray     # [H,W]
    src_points: np.ndarray    # [Ns,3] world (for the splat-rendered GT of dataset.py:362-372)
    src_scales: np.ndarray    # [Ns,3]
    K: np.ndarray             # [3,3]
    W: int
    H: int


def make_scene(rng: np.random.Generator) -> Scene:
    half = np.array([3.0, 1.5, 3.0])
    ns, nb = 6, 4
    sph_r = rng.uniform(0.2, 0.6, ns)
    sph_c = np.stack([rng.uniform(-2.2, 2.2, ns), rng.uniform(-0.9, 0.9, ns), rng.uniform(1.0, 2.4, ns)], 1)
    bc = np.stack([rng.uniform(-2.2, 2.2, nb), rng.uniform(-0.9, 0.9, nb), rng.uniform(1.2, 2.6, nb)], 1)
    bh = rng.uniform(0.15, 0.5, (nb, 3))
    return Scene(half, sph_c, sph_r, bc - bh, bc + bh)


def _rot_xyz(rx: float, ry: float, rz: float) -> np.ndarray:
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def make_pose(rng: np.random.Generator) -> np.ndarray:
    """Camera near the origin looking down +z with a small yaw/pitch."""
    c2w = np.eye(4)
    ang = np.deg2rad(rng.uniform(-10, 10, 2))
    c2w[:3, :3] = _rot_xyz(ang[0], ang[1], 0.0)
    c2w[:3, 3] = rng.uniform(-0.5, 0.5, 3) * np.array([1.0, 0.5, 1.0])
    return c2w


def perturb_pose(c2w: np.ndarray, rng: np.random.Generator, trans_cm: float = 2.0, rot_deg: float = 1.0) -> np.ndarray:
    """pose B = pose A ∘ δ with |δt| ≤ trans_cm per axis and |δR| ≤ rot_deg per axis."""
    d = np.eye(4)
    d[:3, :3] = _rot_xyz(*np.deg2rad(rng.uniform(-rot_deg, rot_deg, 3)))
    d[:3, 3] = rng.uniform(-trans_cm, trans_cm, 3) * 1e-2
    return c2w @ d


def raycast_depth(scene: Scene, c2w: np.ndarray, K: np.ndarray, W: int, H: int) -> np.ndarray:
    """z-depth image; pixel (u,v) shoots the ray K^-1 (u,v,1), u=0..W-1 (kornia depth_to_3d_v2
    convention used by ``src/my_gsplat/geometry.py:158``)."""
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    dc = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], -1).reshape(-1, 3)
    R, o = c2w[:3, :3], c2w[:3, 3]
    d = dc @ R.T                                     # world ray dirs; parameter s == camera z
    best = np.full(d.shape[0], np.inf)
    inv = 1.0 / np.where(np.abs(d) < 1e-12, 1e-12, d)
    # room (hit from inside): exit distance of the slab test
    t1 = (-scene.half - o) * inv
    t2 = (scene.half - o) * inv
    best = np.minimum(best, np.minimum.reduce(np.maximum(t1, t2), axis=1))
    # boxes (hit from outside)
    for lo, hi in zip(scene.box_lo, scene.box_hi):
        a, b = (lo - o) * inv, (hi - o) * inv
        tn = np.maximum.reduce(np.minimum(a, b), axis=1)
        tf = np.minimum.reduce(np.maximum(a, b), axis=1)
        hit = (tn <= tf) & (tn > 1e-3)
        best = np.where(hit & (tn < best), tn, best)
    # spheres
    dd = (d * d).sum(1)
    for c, r in zip(scene.sph_c, scene.sph_r):
        oc = o - c
        bq = d @ oc
        disc = bq * bq - dd * (oc @ oc - r * r)
        s = (-bq - np.sqrt(np.maximum(disc, 0.0))) / dd
        hit = (disc > 0) & (s > 1e-3)
        best = np.where(hit & (s < best), s, best)
    return best.reshape(H, W).astype(np.float32)


def backproject(depth: np.ndarray, K: np.ndarray, c2w: np.ndarray, stride: int) -> np.ndarray:
    """World points of the stride-lattice pixels (``depth_to_points`` + ``transform_points``,
    ``src/my_gsplat/geometry.py:138-161,23-41``)."""
    H, W = depth.shape
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    v, u = np.meshgrid(np.arange(stride // 2, H, stride), np.arange(stride // 2, W, stride), indexing="ij")
    z = depth[v, u].astype(np.float64)
    pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], -1).reshape(-1, 3)
    return (pc @ c2w[:3, :3].T + c2w[:3, 3]).astype(np.float32)


def knn_scales(points: np.ndarray, k: int = 5) -> np.ndarray:
    """``init_gs_scales`` (``src/my_gsplat/geometry.py:44-66``): sqrt(mean(d_i^2)+1e-24) over the
    k-1 nearest neighbours, isotropic.  Host KD-tree, like the reference's small_gicp call."""
    from scipy.spatial import cKDTree

    d, _ = cKDTree(points).query(points, k=k, workers=-1)
    s = np.sqrt((d[:, 1:].astype(np.float64) ** 2).mean(-1) + 1e-24).astype(np.float32)
    return np.repeat(s[:, None], 3, 1)


def make_frame_pair(cam: dict = REPLICA, stride: int = 4, seed: int = 42, trans_cm: float = 2.0,
                    rot_deg: float = 1.0, src_stride: int | None = None, scene: Scene | None = None,
                    tar_c2w: np.ndarray | None = None) -> FramePair:
    rng = np.random.default_rng(seed)
    scene = scene or make_scene(rng)
    W, H = cam["W"], cam["H"]
    K = np.array([[cam["fx"], 0, cam["cx"]], [0, cam["fy"], cam["cy"]], [0, 0, 1]], dtype=np.float64)
    A = make_pose(rng) if tar_c2w is None else tar_c2w.astype(np.float64)
    B = perturb_pose(A, rng, trans_cm, rot_deg)
    dA = raycast_depth(scene, A, K, W, H)
    dB = raycast_depth(scene, B, K, W, H)
    tar = backproject(dA, K, A, stride)
    src = backproject(dB, K, B, src_stride or stride)
    return FramePair(tar, knn_scales(tar), A.astype(np.float32), B.astype(np.float32), dB, src,
                     knn_scales(src), K.astype(np.float32), W, H)


def make_trajectory(n_frames: int, cam: dict = TUM, stride: int = 2, seed: int = 42, trans_cm: float = 6.0,
                    rot_deg: float = 3.0):
    """Smooth random walk of ``n_frames`` poses in one scene; yields FramePair(i, i+1) like
    ``Parser`` iterating a sequence (``src/data/dataset.py:345-383``)."""
    rng = np.random.default_rng(seed)
    scene = make_scene(rng)
    pose = make_pose(rng)
    for i in range(n_frames):
        fp = make_frame_pair(cam, stride, seed=seed * 100003 + i, trans_cm=trans_cm, rot_deg=rot_deg, scene=scene,
                             tar_c2w=pose)
        pose = fp.src_c2w.astype(np.float64)
        # keep the walk inside the room
        pose[:3, 3] = np.clip(pose[:3, 3], [-1.0, -0.5, -1.0], [1.0, 0.5, 1.0])
        yield fp


def random_cloud(N: int, C: int, W: int, H: int, seed: int = 0):
    """Random Gaussians in front of random cameras, the shape of the dependency's own op tests
    (gsplat tests/test_basic.py ``test_isect`` / ``test_projection``)."""
    rng = np.random.default_rng(seed)
    means = (rng.standard_normal((N, 3)) * np.array([1.5, 1.0, 0.8]) + np.array([0, 0, 4.0])).astype(np.float32)
    quats = rng.standard_normal((N, 4)).astype(np.float32)
    scales = np.exp(rng.uniform(-4.0, -1.5, (N, 3))).astype(np.float32)
    opac = rng.uniform(0.05, 1.0, N).astype(np.float32)
    view = np.tile(np.eye(4, dtype=np.float32), (C, 1, 1))
    for c in range(C):
        view[c, :3, :3] = _rot_xyz(*np.deg2rad(rng.uniform(-8, 8, 3)))
        view[c, :3, 3] = rng.uniform(-0.3, 0.3, 3)
    f = 0.9 * W
    Ks = np.tile(np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]], dtype=np.float32), (C, 1, 1))
    return means, quats, scales, opac, view, Ks
