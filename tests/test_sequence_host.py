"""SURVEY.md §8 row f4 on the CPU: dataset readers and the constant-velocity pose bookkeeping against outputs of the
reference's own code (``tests/golden/ref_sequence_small.npz``, made by ``tests/golden/make_golden_sequence.py``), and
the readers end to end on a tiny TUM / Replica directory tree written on the fly."""
import json
import os

import numpy as np
import pytest

from gsplatloc_b200 import datasets as ds
from gsplatloc_b200 import motion

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "ref_sequence_small.npz"))


def test_tum_association_matches_the_reference(gold):
    a3 = ds.associate_frames(gold["t_img"], gold["t_dep"], gold["t_pose"])
    assert np.array_equal(np.array(a3), gold["assoc3"])                       # index work: exact
    assert np.array_equal(np.array(ds.associate_frames(gold["t_img"], gold["t_dep"], None)), gold["assoc2"])
    tight = ds.associate_frames(gold["t_img"], gold["t_dep"], gold["t_pose"], max_dt=0.004)
    assert np.array_equal(np.array(tight), gold["assoc3_tight"])
    for fr in (32, 10, 4):
        assert np.array_equal(np.array(ds.frame_indices(a3, gold["t_img"], fr)), gold[f"indices_fr{fr}"])
    assert ds.associate_frames(np.zeros(0), gold["t_dep"], gold["t_pose"]) == []   # empty lists
    assert ds.frame_indices([], gold["t_img"], 32) == []


def test_tum_pose_vectors_match_the_reference(gold):
    got = np.stack([ds.pose_matrix_from_quaternion(p) for p in gold["pvecs"]])
    np.testing.assert_allclose(got, gold["pose_mats"], rtol=0, atol=1e-14)     # float64 both sides
    rb = ds.rebase_to_first(list(got))
    assert np.array_equal(rb[0], np.eye(4, dtype=np.float32)) and rb[1].dtype == np.float32
    np.testing.assert_allclose(got[0] @ rb[3].astype(np.float64), got[3], atol=1e-5)


def test_rotation_to_quaternion_matches_the_reference_on_every_branch(gold):
    got = np.stack([motion.rotation_matrix_to_quaternion(R) for R in gold["Rs"]])
    np.testing.assert_allclose(got, gold["Rs_quats"], rtol=0, atol=2e-7)       # float32: one or two ulps
    assert (np.abs(gold["Rs_quats"][:, 0]) < 0.1).any()                        # a non-trace branch was exercised


def test_constant_velocity_walk_matches_the_reference(gold):
    walk = gold["walk"]
    m = motion.ConstantVelocity(walk[0])
    q, t, c = [m.quat_cur.copy()], [m.t_cur.copy()], [m.c2w()]
    for i in range(1, 6):
        m.update_pose(walk[i])
        q.append(m.quat_cur.copy()); t.append(m.t_cur.copy()); c.append(m.c2w())
        m.update_pose()
        q.append(m.quat_cur.copy()); t.append(m.t_cur.copy()); c.append(m.c2w())
    np.testing.assert_allclose(np.stack(q), gold["cv_quats"], rtol=0, atol=3e-7)
    np.testing.assert_allclose(np.stack(t), gold["cv_trans"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(np.stack(c), gold["cv_c2w"], rtol=0, atol=2e-6)
    with pytest.raises(ValueError, match="fake new pose"):
        m.update_pose(np.eye(3))


def test_constant_velocity_extrapolates_a_uniform_motion():
    """Property: for a uniform translation the prediction is exact; for a uniform small rotation it is within the
    second-order term of the linear quaternion step."""
    th = np.deg2rad(2.0)

    def pose(i):
        p = np.eye(4)
        c, s = np.cos(i * th), np.sin(i * th)
        p[:3, :3] = [[c, -s, 0], [s, c, 0], [0, 0, 1]]
        p[:3, 3] = [0.03 * i, -0.01 * i, 0.02 * i]
        return p

    m = motion.ConstantVelocity(pose(0))
    m.update_pose(pose(1))
    m.update_pose()
    np.testing.assert_allclose(m.t_cur, pose(2)[:3, 3], atol=1e-6)
    np.testing.assert_allclose(m.c2w()[:3, :3], pose(2)[:3, :3], atol=th * th)


# ------------------------------------------------------------------------------------------------ readers on disk
def _write_depth(path, arr):
    from PIL import Image
    Image.fromarray(arr.astype(np.uint16)).save(path)


def test_tum_reader_on_a_small_tree(tmp_path, gold):
    root = tmp_path / "TUM" / "rgbd_dataset_unit"
    (root / "rgb").mkdir(parents=True); (root / "depth").mkdir()
    cam = {"camera": {"w": 24, "h": 20, "fx": 20.0, "fy": 21.0, "cx": 12.0, "cy": 10.0, "scale": 5000.0, "crop_edge": 2,
                      "distortion": [0.1, 0.0, 0.0, 0.0, 0.0]}}
    (root / "cam_params.json").write_text(json.dumps(cam))
    rng = np.random.default_rng(1)
    t_img, t_dep, t_pose = gold["t_img"][:30], gold["t_dep"][:30], gold["t_pose"][:200]
    imgs = {}
    with open(root / "rgb.txt", "w") as f:
        f.write("# color images\n# file\n# timestamp filename\n")
        for t in t_img:
            f.write(f"{t:.6f} rgb/{t:.6f}.png\n")
    with open(root / "depth.txt", "w") as f:
        f.write("# depth maps\n# file\n# timestamp filename\n")
        for t in t_dep:
            name = f"depth/{t:.6f}.png"
            imgs[name] = rng.integers(0, 40000, (20, 24))
            _write_depth(root / name, imgs[name])
            f.write(f"{t:.6f} {name}\n")
    pv = rng.normal(size=(200, 7))
    with open(root / "groundtruth.txt", "w") as f:
        f.write("# ground truth trajectory\n# file\n# timestamp tx ty tz qx qy qz qw\n")
        for t, p in zip(t_pose, pv):
            f.write(f"{t:.4f} " + " ".join(f"{x:.6f}" for x in p) + "\n")
    seq = ds.TUM("unit", input_folder=tmp_path / "TUM", frame_rate=10)
    assert len(seq) > 2 and len(seq) == len(seq.associations)
    assert seq.K[0, 2] == 10.0 and seq.K[1, 2] == 8.0 and seq.cfg["w"] == 20 and seq.cfg["h"] == 16   # crop_edge
    d, K, c2w = seq[1]
    i, j, k = seq.associations[1]
    name = f"depth/{t_dep[j]:.6f}.png"
    want = (imgs[name].astype(np.float32)[2:-2, 2:-2] / np.float32(5000.0))
    assert d.dtype == np.float32 and d.shape == (16, 20) and np.array_equal(d, want)
    assert np.array_equal(seq.pose(0), np.eye(4, dtype=np.float32)) and c2w.dtype == np.float32
    assert len(seq[0:2]) == 2
    with pytest.raises(ValueError):
        seq[len(seq)]
    with pytest.raises(TypeError):
        seq["0"]


def test_replica_reader_on_a_small_tree(tmp_path):
    root = tmp_path / "Replica"
    res = root / "room0" / "results"
    res.mkdir(parents=True)
    (root / "cam_params.json").write_text(json.dumps(
        {"camera": {"w": 12, "h": 8, "fx": 6.0, "fy": 6.0, "cx": 5.5, "cy": 3.5, "scale": 6553.5}}))
    rng = np.random.default_rng(2)
    depths, poses = [], []
    for i in (0, 1, 2, 10):                                   # natural order: 2 before 10
        depths.append(rng.integers(0, 65535, (8, 12)))
        _write_depth(res / f"depth{i}.png", depths[-1])
        (res / f"frame{i}.jpg").write_bytes(b"")
        poses.append(rng.normal(size=(4, 4)))
    (root / "room0" / "traj.txt").write_text("\n".join(" ".join(f"{x:.17e}" for x in p.ravel()) for p in poses) + "\n")
    seq = ds.Replica("room0", input_folder=root)
    assert len(seq) == 4
    for i in range(4):
        d, K, c2w = seq[i]
        assert np.array_equal(d, (depths[i].astype(np.float64) / 6553.5).astype(np.float32))
        np.testing.assert_allclose(c2w, poses[i], rtol=1e-15)
    # crop_edge on Replica: BaseDataset shrinks h, w, cx, cy (dataset.py:30-35) but Replica._get_depth never crops the
    # image (:116-124; only TUM's does, :215-218) -- the reader reproduces exactly that
    (root / "cam_crop.json").write_text(json.dumps(
        {"camera": {"w": 12, "h": 8, "fx": 6.0, "fy": 6.0, "cx": 5.5, "cy": 3.5, "scale": 6553.5, "crop_edge": 2}}))
    crop = ds.Replica("room0", input_folder=root, cfg_file=root / "cam_crop.json")
    assert crop.cfg["w"] == 8 and crop.cfg["h"] == 4 and crop.K[0, 2] == 3.5 and crop.K[1, 2] == 1.5
    assert crop.depth(0).shape == (8, 12) and np.array_equal(crop.depth(0), seq.depth(0))
    (res / "frame11.jpg").write_bytes(b"")
    with pytest.raises(ValueError, match="do not match"):
        ds.Replica("room0", input_folder=root)
    with pytest.raises(FileNotFoundError):
        ds.Replica("room7", input_folder=root)


# ------------------------------------------------------------------------------------------------ chained walk
class _StandInTracker:
    """Host stand-in for PoseTracker (the chaining logic is host code): every loop "finds" the ground-truth pose plus
    a fixed 1 mm offset along x, and records where it was started."""

    def __init__(self):
        self.inits, self.maps = [], []

    def set_map(self, means, quats, scales, opacities):
        self.maps.append(np.asarray(means))

    def set_observation(self, depth, K):
        pass

    def set_poses(self, c2w_init, c2w_gt=None):
        self.inits.append(np.asarray(c2w_init, dtype=np.float64).copy())
        self._gt = np.asarray(c2w_gt, dtype=np.float64)

    def run(self, n_iters, poll_every=25, use_graph=True):
        return {"iters": 7, "elapsed_ms": 1.0}

    def states(self):
        q = motion.rotation_matrix_to_quaternion(self._gt[:3, :3])
        t = self._gt[:3, 3].astype(np.float32) + np.array([1e-3, 0, 0], dtype=np.float32)
        z = 0.0
        # stopped in iteration 6 (7 executed); the pose ``patience`` steps after the best one has drifted by 5 cm: the
        # chain must hand on the best-loss pose (model.py:41-48), not the last one
        return [dict(iter=6, stopped=1, loss=z, depth_loss=z, edge_loss=z, normal_loss=z, eT=1e-3, eR=z, best_loss=z,
                     best_eT=z, best_eR=z, best_iter=0, q=q, t=t + np.float32(0.05), best_q=q, best_t=t)]


@pytest.mark.parametrize("chain", ["last", "cv"])
def test_track_frames_chains_the_tracked_pose(chain):
    from gsplatloc_b200 import sharding

    def gt(i):                                     # uniform motion: 2 cm per frame along z
        p = np.eye(4)
        p[2, 3] = 0.02 * i
        return p

    seen = []

    def load_frame(f, c2w_map):
        seen.append(None if c2w_map is None else np.asarray(c2w_map, dtype=np.float64).copy())
        return dict(means=np.zeros((1, 3)), quats=None, scales=None, opacities=None, depth=None, K=np.eye(3),
                    c2w_init=gt(f), c2w_gt=gt(f + 1))

    trk = _StandInTracker()
    res = sharding.track_frames([3, 4, 5, 6], load_frame, trk, rank=0, chain=chain)
    assert res.iters == 28 and res.records.shape == (4, sharding.RECORD_WIDTH)
    assert seen[0] is None                                       # the chain starts at the first frame's own pose
    off = np.array([1e-3, 0, 0])
    for n, f in enumerate([4, 5, 6], start=1):                   # later maps sit at the TRACKED pose of the loop before
        np.testing.assert_allclose(seen[n][:3, 3], gt(f)[:3, 3] + off, atol=1e-6)
    np.testing.assert_allclose(trk.inits[0], gt(3), atol=1e-7)
    if chain == "last":
        np.testing.assert_allclose(trk.inits[2][:3, 3], gt(5)[:3, 3] + off, atol=1e-6)
    else:                                                        # uniform motion is extrapolated exactly from loop 3 on
        np.testing.assert_allclose(trk.inits[2][:3, 3], gt(6)[:3, 3] + off, atol=1e-6)
        np.testing.assert_allclose(trk.inits[3][:3, 3], gt(7)[:3, 3] + off, atol=1e-6)
    with pytest.raises(ValueError):
        sharding.track_frames([0], load_frame, trk, chain="slerp")
