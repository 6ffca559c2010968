#!/usr/bin/env python
"""Generates ``tests/golden/ref_geometry_small.npz`` by running GsplatLoc's OWN per-frame setup code
(``/root/reference/src/my_gsplat/geometry.py``: ``init_gs_scales``, ``transform_points``, ``depth_to_points``,
``compute_depth_gt``; ``src/my_gsplat/utils.py``: ``knn``) on small seeded inputs, in the build container (CPU):

    python tests/golden/make_golden_geometry.py

Shims (``ref_shims.py``): ``small_gicp.KdTree`` is scipy's cKDTree returning squared distances, kornia's
``depth_to_3d_v2`` is restated, ``gsplat.rasterization`` is the CPU oracle.  So the fixture pins the reference's own
arithmetic around those three (the double squaring of init_gs_scales, eps, the repeat; addmm; the flattening order of
depth_to_points; the argument set compute_depth_gt hands to rasterization) and is a regression pin for the rest.
"""
from __future__ import annotations

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch

import gsb_testutil as U
import ref_shims
from oracle import gsplat_oracle as orc


def main():
    torch.manual_seed(42)
    ref = ref_shims.install(orc.rasterization)
    geo = ref["geometry"]
    rng = np.random.default_rng(11)
    out = {}
    # (1) a depth-lattice cloud (what the loop sees) and (2) a volumetric cloud with duplicates and an outlier
    fp = U.frame(U.SMALL_CAM, stride=2, seed=42)
    lattice = fp.tar_points.astype(np.float32)
    blob = rng.normal(size=(3000, 3)).astype(np.float32) * np.array([1.0, 0.3, 2.0], dtype=np.float32)
    blob[10] = blob[11]                      # duplicate point: a zero neighbour distance
    blob[20] = np.array([40.0, -35.0, 60.0])  # far outlier: many empty grid shells
    for tag, pts in (("lattice", lattice), ("blob", blob)):
        t = torch.tensor(pts)
        out[tag + "_points"] = pts
        for k in (5, 4):
            out[f"{tag}_knn{k}"] = ref["geometry"].knn(t, k).detach().numpy().astype(np.float32)
            out[f"{tag}_scales{k}"] = geo.init_gs_scales(t, k).detach().numpy().astype(np.float32)
    # transform_points / depth_to_points
    T = fp.src_c2w.astype(np.float32)
    out["T"] = T
    out["lattice_transformed"] = geo.transform_points(torch.tensor(T), torch.tensor(lattice)).numpy()
    depth = fp.src_depth.astype(np.float32)
    out["depth"] = depth
    out["K"] = fp.K.astype(np.float32)
    out["depth_points"] = geo.depth_to_points(torch.tensor(depth), torch.tensor(fp.K.astype(np.float32))).numpy()
    # compute_depth_gt: the source cloud re-rendered from the target pose (dataset.py:362-372)
    src = fp.src_points.astype(np.float32) if hasattr(fp, "src_points") else None
    if src is None:
        src = geo.transform_points(torch.tensor(T), geo.depth_to_points(torch.tensor(depth), torch.tensor(out["K"]))).numpy()
        src = src[:: 4]
    out["gt_points"] = src
    out["gt_c2w"] = fp.tar_c2w.astype(np.float32)
    gt = geo.compute_depth_gt(torch.tensor(src), torch.zeros(src.shape[0], 3), torch.tensor(out["K"])[None],
                              torch.tensor(out["gt_c2w"])[None], fp.H, fp.W)
    out["gt_depth"] = gt.detach().numpy().astype(np.float32)
    # PCA normalisation (src/data/normalize.py): the reference's own align_principle_axes / transform_cameras
    import importlib
    nz = importlib.import_module("src.data.normalize")
    for tag, pts in (("lattice", lattice), ("blob", blob)):
        out[tag + "_pca_T"] = nz.align_principle_axes(torch.tensor(pts)).numpy()
    poses = torch.tensor(np.stack([fp.tar_c2w, fp.src_c2w]).astype(np.float32))
    tc, sc = nz.transform_cameras(torch.tensor(out["lattice_pca_T"]), poses.clone())
    out["pca_poses_in"] = poses.numpy(); out["pca_poses_out"] = tc.numpy(); out["pca_scaling"] = sc.numpy()
    out["cam"] = np.array([fp.W, fp.H], dtype=np.int64)
    path = os.path.join(HERE, "ref_geometry_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
    for k_, v in out.items():
        print(k_, v.shape, v.dtype)


if __name__ == "__main__":
    main()
