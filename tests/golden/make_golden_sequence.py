#!/usr/bin/env python
"""Generates ``tests/golden/ref_sequence_small.npz`` by running GsplatLoc's OWN sequence bookkeeping in the build
container (CPU): the TUM association helpers (``/root/reference/src/data/dataset.py:279-321``, static methods of
``TUM``) on seeded timestamp lists, and ``CameraOptModule_quat_tans`` (``src/my_gsplat/model.py:27-82``) driven through
``update_pose`` / ``predict_next_pose`` / ``forward`` over a short walk.

    python tests/golden/make_golden_sequence.py

Shims (``ref_shims.py``): kornia's quaternion functions are restated; everything else here is the reference's code.
"""
from __future__ import annotations

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import importlib

import numpy as np
import torch

import ref_shims
from oracle import gsplat_oracle as orc


def rot(rng, deg):
    a = rng.normal(size=3)
    a /= np.linalg.norm(a)
    th = np.deg2rad(deg)
    Kx = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx


def main():
    ref = ref_shims.install(orc.rasterization)
    TUM = importlib.import_module("src.data.dataset").TUM
    rng = np.random.default_rng(5)
    out = {}
    # ---- TUM lists: 30 Hz colour, depth with jitter and a dropped stretch, 100 Hz poses starting late
    t_img = 1305031100.0 + np.arange(120) / 30.0 + rng.normal(0, 2e-3, 120)
    t_dep = np.delete(1305031100.01 + np.arange(120) / 30.0 + rng.normal(0, 4e-3, 120), np.s_[40:47])
    t_pose = 1305031100.6 + np.arange(500) / 100.0
    out["t_img"], out["t_dep"], out["t_pose"] = t_img, t_dep, t_pose
    assoc = TUM._associate_frames(t_img, t_dep, t_pose)
    out["assoc3"] = np.array(assoc, dtype=np.int64)
    out["assoc2"] = np.array(TUM._associate_frames(t_img, t_dep, None), dtype=np.int64)
    out["assoc3_tight"] = np.array(TUM._associate_frames(t_img, t_dep, t_pose, max_dt=0.004), dtype=np.int64)
    for fr in (32, 10, 4):
        out[f"indices_fr{fr}"] = np.array(TUM._get_frame_indices(assoc, t_img, fr), dtype=np.int64)
    pv = rng.normal(size=(16, 7))
    pv[:, 3:] *= rng.uniform(0.5, 2.0, (16, 1))          # un-normalised quaternions: from_quat normalises
    out["pvecs"] = pv
    out["pose_mats"] = np.stack([TUM._pose_matrix_from_quaternion(p) for p in pv])

    # ---- CameraOptModule_quat_tans over a walk: best pose in, constant-velocity prediction out
    Mod = ref["model"].CameraOptModule_quat_tans
    poses = []
    c2w = np.eye(4)
    c2w[:3, :3] = rot(rng, 170.0)                         # near-pi rotation: a non-trace branch of the conversion
    c2w[:3, 3] = rng.normal(size=3)
    for i in range(6):
        poses.append(c2w.copy())
        step = np.eye(4)
        step[:3, :3] = rot(rng, 3.0 + i)
        step[:3, 3] = rng.normal(size=3) * 0.05
        c2w = c2w @ step
    poses = np.stack(poses).astype(np.float32)
    out["walk"] = poses
    m = Mod(torch.tensor(poses[0]))
    q, t, fwd = [], [], []
    with torch.no_grad():
        q.append(m.quat_cur.detach().numpy().copy()); t.append(m.t_cur.detach().numpy().copy())
        fwd.append(m.forward().numpy().copy())
        for i in range(1, 6):
            m.update_pose(torch.tensor(poses[i]))         # the loop's best pose
            q.append(m.quat_cur.detach().numpy().copy()); t.append(m.t_cur.detach().numpy().copy())
            fwd.append(m.forward().numpy().copy())
            m.update_pose()                               # constant-velocity prediction for the next loop
            q.append(m.quat_cur.detach().numpy().copy()); t.append(m.t_cur.detach().numpy().copy())
            fwd.append(m.forward().numpy().copy())
    out["cv_quats"], out["cv_trans"], out["cv_c2w"] = np.stack(q), np.stack(t), np.stack(fwd)
    # rotation_matrix_to_quaternion over all four branches
    Rs = np.stack([rot(rng, d) for d in (1.0, 60.0, 120.0, 179.0, 179.5, 178.0, 175.0, 90.0)]).astype(np.float32)
    out["Rs"] = Rs
    out["Rs_quats"] = ref["transform"].rotation_matrix_to_quaternion(torch.tensor(Rs)).numpy()
    path = os.path.join(HERE, "ref_sequence_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
