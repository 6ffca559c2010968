"""Host side of chained tracking: the pose bookkeeping of ``CameraOptModule_quat_tans``
(``src/my_gsplat/model.py:27-82``) that sits BETWEEN two pose loops — where the next loop starts from.  The loop
itself (quaternion + translation Adam on the device) is ``PoseTracker``; this class only keeps the two last poses
as (w, x, y, z) quaternion + translation in float32, like the reference's parameters, and hands
``PoseTracker.set_poses`` its initial pose.  Seven floats of host arithmetic per frame: no kernel.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .tracker import quat_t_to_c2w

_F = np.float32


def rotation_matrix_to_quaternion(R: np.ndarray, eps: float = 1e-8) -> np.ndarray:
    """``rotation_matrix_to_quaternion`` (transform.py:84 over kornia 0.7.2): trace method, (w, x, y, z), float32,
    ``eps`` under each square root, branch order trace > 0, then m00 largest, then m11 > m22."""
    m = np.asarray(R, dtype=_F).reshape(3, 3)
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = (m[i, j] for i in range(3) for j in range(3))
    one, e, tiny = _F(1.0), _F(eps), np.finfo(_F).tiny

    def div(n, d):
        return _F(n) / max(_F(d), tiny)

    if m00 + m11 + m22 > 0:
        s = np.sqrt(m00 + m11 + m22 + one + e) * _F(2.0)
        q = (_F(0.25) * s, div(m21 - m12, s), div(m02 - m20, s), div(m10 - m01, s))
    elif m00 > m11 and m00 > m22:
        s = np.sqrt(one + m00 - m11 - m22 + e) * _F(2.0)
        q = (div(m21 - m12, s), _F(0.25) * s, div(m01 + m10, s), div(m02 + m20, s))
    elif m11 > m22:
        s = np.sqrt(one + m11 - m00 - m22 + e) * _F(2.0)
        q = (div(m02 - m20, s), div(m01 + m10, s), _F(0.25) * s, div(m12 + m21, s))
    else:
        s = np.sqrt(one + m22 - m00 - m11 + e) * _F(2.0)
        q = (div(m10 - m01, s), div(m02 + m20, s), div(m12 + m21, s), _F(0.25) * s)
    return np.array(q, dtype=_F)


def normalize_quaternion(q: np.ndarray, eps: float = 1e-12) -> np.ndarray:
    """kornia ``normalize_quaternion``: q / max(|q|, eps), float32."""
    q = np.asarray(q, dtype=_F)
    return (q / max(_F(np.sqrt(np.sum(q * q, dtype=_F))), _F(eps))).astype(_F)


class ConstantVelocity:
    """Pose state between loops, named after what ``update_pose`` / ``predict_next_pose`` (model.py:41-77) do.

    ``update_pose(c2w)``   the loop's best pose becomes the current one (the reference also re-creates its Adam
                           optimisers there: ``PoseTracker.set_poses`` resets the moments the same way);
    ``update_pose()``      the current pose is replaced by the constant-velocity prediction: quaternion and translation
                           each advanced by their last difference, the quaternion re-normalised (a linear step in
                           quaternion space, as the reference does, not a slerp);
    ``c2w()``              ``forward()``: the current pose as a 4x4 camera-to-world matrix.
    """

    def __init__(self, init_pose: np.ndarray):
        p = np.asarray(init_pose, dtype=_F).reshape(4, 4)
        self.quat_cur = rotation_matrix_to_quaternion(p[:3, :3])
        self.t_cur = p[:3, 3].copy()
        self.prev_quat = self.quat_cur.copy()
        self.prev_t = self.t_cur.copy()

    def update_pose(self, new_pose: Optional[np.ndarray] = None) -> None:
        if new_pose is None:
            self.quat_cur, self.t_cur = self.predict_next_pose()
            return
        p = np.asarray(new_pose)
        if p.shape != (4, 4):
            raise ValueError("fake new pose")                       # model.py:57-58
        p = p.astype(_F)
        self.quat_cur = rotation_matrix_to_quaternion(p[:3, :3])
        self.t_cur = p[:3, 3].copy()

    def predict_next_pose(self) -> Tuple[np.ndarray, np.ndarray]:
        q = normalize_quaternion(self.quat_cur + (self.quat_cur - self.prev_quat))
        t = (self.t_cur + (self.t_cur - self.prev_t)).astype(_F)
        self.prev_quat, self.prev_t = self.quat_cur.copy(), self.t_cur.copy()
        return q, t

    def c2w(self) -> np.ndarray:
        return quat_t_to_c2w(self.quat_cur, self.t_cur).astype(_F)
