"""Typed counterparts of the reference's per-frame records (``src/data/base.py:81-125``): what ``Parser.__getitem__``
(``src/data/dataset.py:345-383``) hands to ``Runner.train`` (``gs_trainer_total.py:49-63``) for one frame pair.

Same field names and shapes, so that code written against the reference's ``AlignData`` reads the same.  Differences,
on purpose: ``__post_init__`` does not move the tensors to a global ``DEVICE`` nor switch ``requires_grad`` on for every
field (the loop reads no gradient of them); the colour half (``colors``, ``pixels``) is optional
because the tracked path is depth only.
"""
from __future__ import annotations

from dataclasses import dataclass, fields
from typing import Optional

import numpy as np
import torch
from torch import Tensor


@dataclass
class TensorWrapper:
    """``base.py:81-106``: ``to(device)`` moves every tensor attribute."""

    def to(self, device) -> "TensorWrapper":
        for f in fields(self):
            v = getattr(self, f.name)
            if torch.is_tensor(v):
                setattr(self, f.name, v.to(device))
        return self


@dataclass
class AlignData(TensorWrapper):
    """``base.py:109-125``: one (target, source) frame pair in the PCA-normalised frame."""
    tar_points: Tensor                 # [N, 3]  map of the loop (back-projected target frame)
    src_points: Tensor                 # [M, 3]
    src_depth: Tensor                  # [1, H, W, 1]  observation (compute_depth_gt re-render of the source frame)
    tar_c2w: Tensor                    # [4, 4]  start pose of the loop
    src_c2w: Tensor                    # [4, 4]  ground truth of the loop
    tar_nums: int                      # len(tar_points)
    pca_factor: Tensor                 # scalar the normalisation divided by (normalize.py:103-124)
    colors: Optional[Tensor] = None    # [N, 3]      (unused on the depth path)
    pixels: Optional[Tensor] = None    # [1, H, W, 3] (unused on the depth path)

    def as_frame(self, K, scales: Optional[Tensor] = None, knn: int = 5) -> dict:
        """The dict ``sharding.track_frames`` consumes: map = ``GSModel(tar_points)`` (model.py:136-178: identity
        quaternions, opacity sigmoid(logit(1)) = 1, isotropic KNN scales from ``geometry.init_gs_scales``)."""
        from . import geometry as geo
        pts = self.tar_points.contiguous()
        n = pts.shape[0]
        quats = torch.zeros((n, 4), dtype=pts.dtype, device=pts.device)
        quats[:, 0] = 1.0
        if scales is None:
            scales = geo.init_gs_scales(pts, knn)
        return dict(means=pts, quats=quats, scales=scales, opacities=torch.ones(n, dtype=pts.dtype, device=pts.device),
                    depth=self.src_depth.reshape(self.src_depth.shape[-3], self.src_depth.shape[-2]).contiguous(),
                    K=np.asarray(K.cpu() if torch.is_tensor(K) else K, dtype=np.float32),
                    c2w_init=np.asarray(self.tar_c2w.cpu(), dtype=np.float32), c2w_gt=np.asarray(self.src_c2w.cpu(), dtype=np.float32))
