"""Dataset readers either side of the pose loop: the depth + pose half of the reference's
``Replica`` and ``TUM`` classes (``src/data/dataset.py:17-321``) — the colour images never reach this path
(``use_rgb`` stays off in the shipped trainers), so they are listed but not decoded.

Host logic only: file lists, the TUM timestamp association, pose parsing, PNG depth / scale.  A sequence item is
``(depth [H, W] float32 metres, K [3, 3], c2w [4, 4])``; hand the depth to ``geometry.depth_to_points`` and the rest
of the per-frame setup runs on the GPU (``tools/track_sequence.py``).
"""
from __future__ import annotations

import json
import re
from pathlib import Path
from typing import List, Optional, Sequence, Tuple

import numpy as np


def load_camera_cfg(cfg_path) -> dict:
    """utils.py:13-27: YAML or JSON camera file; the ``camera`` entry holds w, h, fx, fy, cx, cy, scale
    (+ distortion, crop_edge)."""
    p = Path(cfg_path)
    if not p.exists():
        raise FileNotFoundError(f"File not found: {p}")
    with open(p) as f:
        if p.suffix in (".yaml", ".yml"):
            import yaml
            return yaml.safe_load(f)
        if p.suffix == ".json":
            return json.load(f)
    raise TypeError(f"Failed to load cfg via:{p.suffix}")


def as_intrinsics_matrix(intrinsics: Sequence[float]) -> np.ndarray:
    """utils.py:46-57: [fx, fy, cx, cy] -> K."""
    K = np.eye(3)
    K[0, 0], K[1, 1], K[0, 2], K[1, 2] = intrinsics
    return K


def natural_key(path) -> list:
    """Ordering of ``natsorted`` for the frame files: digit runs compare as numbers (frame2 < frame10)."""
    return [int(s) if s.isdigit() else s for s in re.split(r"(\d+)", Path(path).as_posix())]


def read_depth_png(path) -> np.ndarray:
    """``cv2.imread(path, IMREAD_UNCHANGED)`` for the 16-bit single-channel depth PNGs of both datasets."""
    from PIL import Image
    with Image.open(path) as im:
        a = np.array(im)
    if a.ndim != 2:
        raise ValueError(f"{path}: depth image with {a.ndim} dimensions, expected a single channel")
    return a


# --------------------------------------------------------------------------------------------- TUM association
def associate_frames(tstamp_image: np.ndarray, tstamp_depth: np.ndarray, tstamp_pose: Optional[np.ndarray],
                     max_dt: float = 0.08) -> List[tuple]:
    """dataset.py:279-301: for every colour timestamp the nearest depth (and pose) timestamp, kept when both are closer
    than ``max_dt`` seconds.  Vectorised (one |a - b| table per list), same first-minimum tie rule as ``np.argmin``."""
    ti = np.asarray(tstamp_image, dtype=np.float64)
    td = np.asarray(tstamp_depth, dtype=np.float64)
    if ti.size == 0 or td.size == 0:
        return []
    dd = np.abs(td[None, :] - ti[:, None])
    j = dd.argmin(axis=1)
    ok = dd[np.arange(ti.size), j] < max_dt
    if tstamp_pose is None:
        return [(int(i), int(j[i])) for i in np.nonzero(ok)[0]]
    tp = np.asarray(tstamp_pose, dtype=np.float64)
    if tp.size == 0:
        return []
    dp = np.abs(tp[None, :] - ti[:, None])
    k = dp.argmin(axis=1)
    ok &= dp[np.arange(ti.size), k] < max_dt
    return [(int(i), int(j[i]), int(k[i])) for i in np.nonzero(ok)[0]]


def frame_indices(associations: Sequence[tuple], tstamp_image: np.ndarray, frame_rate: float) -> List[int]:
    """dataset.py:303-314: keep an association when it lies more than 1 / frame_rate after the last kept one."""
    if len(associations) == 0:
        return []
    ti = np.asarray(tstamp_image, dtype=np.float64)
    keep = [0]
    for i in range(1, len(associations)):
        if ti[associations[i][0]] - ti[associations[keep[-1]][0]] > 1.0 / frame_rate:
            keep.append(i)
    return keep


def pose_matrix_from_quaternion(pvec: Sequence[float]) -> np.ndarray:
    """dataset.py:316-321: [tx, ty, tz, qx, qy, qz, qw] (TUM ground-truth order) -> c2w; the quaternion is normalised
    first, as scipy's ``Rotation.from_quat`` does."""
    v = np.asarray(pvec, dtype=np.float64)
    x, y, z, w = v[3:7] / np.linalg.norm(v[3:7])
    pose = np.eye(4)
    pose[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                    [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                    [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    pose[:3, 3] = v[:3]
    return pose


def rebase_to_first(poses: Sequence[np.ndarray]) -> List[np.ndarray]:
    """dataset.py:259-270: TUM poses are expressed relative to the first kept frame (which becomes the identity)."""
    out, inv0 = [], None
    for c2w in poses:
        if inv0 is None:
            inv0 = np.linalg.inv(c2w)
            c2w = np.eye(4)
        else:
            c2w = inv0 @ c2w
        out.append(c2w.astype(np.float32))
    return out


def _parse_list(path: Path, skiprows: int = 0) -> np.ndarray:
    """dataset.py:274-276: space-separated text table, '#' comment lines dropped, everything kept as strings."""
    return np.loadtxt(path, delimiter=" ", dtype=np.str_, skiprows=skiprows, ndmin=2)


# --------------------------------------------------------------------------------------------- sequences
class DepthSequence:
    """``BaseDataset`` (dataset.py:17-76) without the colour half: camera file, crop, K, indexing rules."""

    def __init__(self, input_folder, cfg_file):
        self.input_folder = Path(input_folder)
        if not self.input_folder.exists():
            raise FileNotFoundError(f"Path {input_folder} does not exist.")
        self.cfg = dict(load_camera_cfg(cfg_file)["camera"])
        self.scale = self.cfg["scale"]
        self.distortion = np.array(self.cfg["distortion"]) if "distortion" in self.cfg else None
        self.crop_edge = self.cfg.get("crop_edge", 0)
        if self.crop_edge:
            self.cfg["h"] -= 2 * self.crop_edge
            self.cfg["w"] -= 2 * self.crop_edge
            self.cfg["cx"] -= self.crop_edge
            self.cfg["cy"] -= self.crop_edge
        self.K = as_intrinsics_matrix([self.cfg["fx"], self.cfg["fy"], self.cfg["cx"], self.cfg["cy"]])
        self._depth_paths: List[Path] = []
        self._color_paths: List[Path] = []
        self._poses: List[np.ndarray] = []

    def __len__(self) -> int:
        return len(self._depth_paths)

    _divide_in = np.float32          # TUM divides in float32 (dataset.py:204-213), Replica in float64 (:116-124)
    _crops_depth = True              # only TUM._get_depth crops the image (dataset.py:215-218); Replica's never does,
                                     # although BaseDataset shrinks h, w, cx, cy for both (:30-35)

    def depth(self, index: int) -> np.ndarray:
        d = read_depth_png(self._depth_paths[index]).astype(self._divide_in)
        if self.crop_edge > 0 and self._crops_depth:
            d = d[self.crop_edge:-self.crop_edge, self.crop_edge:-self.crop_edge]
        return (d / self.scale).astype(np.float32)   # float32 is what to_tensor makes of it before the loop

    def pose(self, index: int) -> np.ndarray:
        return self._poses[index]

    def _get_one(self, index: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        return self.depth(index), self.K, self.pose(index)

    def __getitem__(self, index):
        if isinstance(index, (int, np.integer)):
            if index >= len(self) or index < 0:
                raise ValueError(f"Index {index} out of range (0 to {len(self) - 1})")
            return self._get_one(int(index))
        if isinstance(index, slice):
            return [self._get_one(i) for i in range(*index.indices(len(self)))]
        raise TypeError(f"index must be int or slice but now is {type(index)}")


class Replica(DepthSequence):
    """dataset.py:78-161: ``<folder>/<name>/results/{frame*.jpg, depth*.png}`` in natural order, ``traj.txt`` with one
    row-major 4x4 c2w per line, camera file shared by all scenes."""
    _divide_in = np.float64
    _crops_depth = False

    def __init__(self, name: str = "room0", *, input_folder="datasets/Replica", cfg_file=None):
        root = Path(input_folder)
        super().__init__(root / name, cfg_file if cfg_file is not None else root / "cam_params.json")
        self.name = name
        self._color_paths = sorted(self.input_folder.rglob("frame*.jpg"), key=natural_key)
        self._depth_paths = sorted(self.input_folder.rglob("depth*.png"), key=natural_key)
        if len(self._color_paths) == 0 or len(self._depth_paths) == 0:
            raise FileNotFoundError(f"No images found in {self.input_folder}. Please check the path.")
        if len(self._color_paths) != len(self._depth_paths):
            raise ValueError(f"Number of color and depth images do not match in {self.input_folder}.")
        with open(self.input_folder / "traj.txt") as f:
            lines = f.readlines()
        self._poses = [np.array(list(map(float, lines[i].split()))).reshape(4, 4) for i in range(len(self))]


class TUM(DepthSequence):
    """dataset.py:164-321: ``rgbd_dataset_<name>/`` with rgb.txt, depth.txt, groundtruth.txt (or pose.txt) and its own
    cam_params.json; frames associated by timestamp (80 ms), thinned to ``frame_rate``, poses rebased to the first."""

    def __init__(self, name: str = "freiburg1_desk", *, input_folder="datasets/TUM", frame_rate: float = 32):
        self.name = "rgbd_dataset_" + name
        data_dir = Path(input_folder) / self.name
        super().__init__(data_dir, data_dir / "cam_params.json")
        d = self.input_folder
        pose_list = d / ("groundtruth.txt" if (d / "groundtruth.txt").is_file() else "pose.txt")
        image_data = _parse_list(d / "rgb.txt")
        depth_data = _parse_list(d / "depth.txt")
        pose_data = _parse_list(pose_list, skiprows=1)
        pose_vecs = pose_data[:, 1:].astype(np.float64)
        t_img = image_data[:, 0].astype(np.float64)
        assoc = associate_frames(t_img, depth_data[:, 0].astype(np.float64), pose_data[:, 0].astype(np.float64))
        self.associations = [assoc[ix] for ix in frame_indices(assoc, t_img, frame_rate)]
        self._color_paths = [d / image_data[i, 1] for i, _, _ in self.associations]
        self._depth_paths = [d / depth_data[j, 1] for _, j, _ in self.associations]
        self._poses = rebase_to_first([pose_matrix_from_quaternion(pose_vecs[k]) for _, _, k in self.associations])
