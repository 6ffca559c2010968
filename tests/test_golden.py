"""The CPU oracle against the committed golden fixture ``tests/golden/ref_loop_small.npz`` — produced by running
GsplatLoc's own loop code (CameraOptModule_quat_tans, GSModel.forward, loss.py, geometry.py, eval/utils.py) with
``tests/golden/make_golden.py``; see that script for what is and is not pinned (gsplat's rasteriser is not: the
submodule is absent, the fixture's renderer is the oracle itself).  Tolerances: values 1e-5 relative (fp32 fixture vs
fp64 oracle), gradients 1e-4 relative."""
import os

import numpy as np
import pytest
import torch

from oracle import gsplat_oracle as orc
import gsb_testutil as U

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_loop_small.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLDEN))


def _scene(gold, dt):
    N = gold["means"].shape[0]
    means = torch.tensor(gold["means"], dtype=dt)
    quats = torch.tensor([[1.0, 0, 0, 0]], dtype=dt).repeat(N, 1)
    scales = torch.tensor(gold["scales"], dtype=dt)
    W, H = int(gold["cam"][0]), int(gold["cam"][1])
    return means, quats, scales, torch.ones(N, dtype=dt), torch.tensor(gold["gt_depth"], dtype=dt), \
        torch.tensor(gold["K"], dtype=dt), W, H


def test_fixture_inputs_are_the_seeded_frame(gold):
    fp = U.frame(U.SMALL_CAM, stride=int(gold["frame_args"][0]), seed=int(gold["frame_args"][1]))
    assert np.array_equal(fp.tar_points.astype(np.float32), gold["means"])
    assert np.array_equal(fp.src_depth.astype(np.float32), gold["gt_depth"])


@pytest.mark.parametrize("tag,dl,nl", [("A", 0.8, 0.0), ("B", 0.7, 0.1)])
def test_oracle_step_matches_reference_loop_step0(gold, tag, dl, nl):
    dt = torch.float64
    means, quats, scales, opac, gt, K, W, H = _scene(gold, dt)
    q = torch.tensor(gold[tag + "_quat"][0], dtype=dt).requires_grad_(True)
    t = torch.tensor(gold[tag + "_t"][0], dtype=dt).requires_grad_(True)
    tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q, t, means, quats, scales, opac, gt, K, W, H, dl, nl, 1)
    d.retain_grad()
    tot.backward()
    assert U.rel_err(orc.pose_to_c2w(q.detach(), t.detach()).numpy(), gold[tag + "_c2w"][0]) < 1e-6
    assert U.rel_err(d.detach().numpy(), gold[tag + "_first_depth"]) < 1e-5
    assert U.rel_err(a.detach().numpy(), gold[tag + "_first_alpha"]) < 1e-5
    assert abs(float(Ld) - gold[tag + "_Ld"][0]) / gold[tag + "_Ld"][0] < 1e-5
    assert abs(float(Ls) - gold[tag + "_Ls"][0]) / gold[tag + "_Ls"][0] < 1e-5
    if nl:
        assert abs(float(Ln) - gold[tag + "_Ln"][0]) / gold[tag + "_Ln"][0] < 1e-5
    assert abs(float(tot) - gold[tag + "_total"][0]) / gold[tag + "_total"][0] < 1e-5
    assert U.rel_err(q.grad.numpy(), gold[tag + "_g_quat"][0]) < 1e-4
    assert U.rel_err(t.grad.numpy(), gold[tag + "_g_t"][0]) < 1e-4
    eT, eR = orc.pose_errors(gold[tag + "_c2w"][0], gold["c2w_gt"])
    assert abs(eT - gold[tag + "_eT"][0]) < 1e-6 and abs(eR - gold[tag + "_eR"][0]) < 2e-3


def test_oracle_loop_follows_reference_adam_trajectory(gold):
    """Adam with coupled weight decay + ExponentialLR stepped every iteration: 6 steps of the reference's loop."""
    dt = torch.float32
    means, quats, scales, opac, gt, K, W, H = _scene(gold, dt)
    n = gold["A_quat"].shape[0]
    grads = []
    res = orc.pose_opt_loop(torch.tensor(gold["A_quat"][0]), torch.tensor(gold["A_t"][0]), means, quats, scales, opac,
                            gt, K, W, H, c2w_gt=gold["c2w_gt"], max_steps=int(gold["max_steps"]), n_steps=n,
                            early_stop=False, log=lambda s, row, gq, gt_: grads.append((gq.numpy(), gt_.numpy())))
    hist = np.array(res["history"])
    assert np.allclose(hist[:, 0], gold["A_total"], rtol=2e-5)
    assert np.allclose(hist[:, 3], gold["A_eT"], atol=2e-6)
    for s in range(n):
        assert U.rel_err(grads[s][0], gold["A_g_quat"][s]) < 1e-3
        assert U.rel_err(grads[s][1], gold["A_g_t"][s]) < 1e-3
    assert np.allclose(res["quat"].numpy(), gold["A_final_quat"], atol=2e-6)
    assert np.allclose(res["t"].numpy(), gold["A_final_t"], atol=2e-6)
    gamma = 0.2 ** (1.0 / int(gold["max_steps"]))
    assert np.allclose(gold["A_lr_q"], 5e-4 * gamma ** np.arange(n), rtol=1e-6)
    assert np.allclose(gold["A_lr_t"], 1e-3 * gamma ** np.arange(n), rtol=1e-6)


def test_pose_module_and_metrics_match_reference(gold):
    for pose, quat, c2w, other, eT, eR in zip(gold["P_pose"], gold["P_quat"], gold["P_c2w"], gold["P_other"],
                                               gold["P_eT"], gold["P_eR"]):
        q0 = orc.rotmat_to_quat(pose[:3, :3])
        scale = np.linalg.norm(quat)                       # the fixture rescales quat_cur after initialisation
        assert np.allclose(q0 * scale, quat, atol=2e-6)
        c = orc.pose_to_c2w(torch.tensor(quat, dtype=torch.float64), torch.tensor(pose[:3, 3], dtype=torch.float64))
        assert np.allclose(c.numpy(), c2w, atol=2e-6)
        e = orc.pose_errors(c2w, other)
        assert abs(e[0] - eT) < 1e-6 and abs(e[1] - eR) < 5e-3


# ------------------------------------------------------------------------------- per-frame setup (rows f1, f2)
GOLDEN_GEO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_geometry_small.npz")


@pytest.fixture(scope="module")
def gold_geo():
    return dict(np.load(GOLDEN_GEO))


@pytest.mark.parametrize("tag", ["lattice", "blob"])
@pytest.mark.parametrize("k", [5, 4])
def test_geometry_oracle_knn_and_scales_match_reference_code(gold_geo, tag, k):
    """oracle/geometry_oracle.py against the reference's own knn + init_gs_scales (make_golden_geometry.py).
    Tolerance: 1e-6 relative (float32 fixture)."""
    from oracle import geometry_oracle as geo
    pts = gold_geo[tag + "_points"]
    d = geo.knn_sq_dists(pts, k).astype(np.float32)
    assert np.allclose(d, gold_geo[f"{tag}_knn{k}"], rtol=1e-6, atol=1e-12)
    s = geo.init_gs_scales(pts, k)
    assert s.shape == (pts.shape[0], 3)
    assert np.allclose(s, gold_geo[f"{tag}_scales{k}"], rtol=2e-6, atol=0)
    # the non-literal variant is the RMS distance to the k-1 neighbours
    rms = geo.init_gs_scales(pts, k, reference_faithful=False)[:, 0]
    assert np.allclose(rms, np.sqrt(gold_geo[f"{tag}_knn{k}"][:, 1:].mean(-1) + 1e-24), rtol=2e-6)


def test_geometry_oracle_points_match_reference_code(gold_geo):
    from oracle import geometry_oracle as geo
    assert np.allclose(geo.transform_points(gold_geo["T"], gold_geo["lattice_points"]), gold_geo["lattice_transformed"],
                       rtol=1e-6, atol=1e-6)
    p = geo.depth_to_points(gold_geo["depth"], gold_geo["K"])
    assert p.shape == gold_geo["depth_points"].shape
    assert np.array_equal(p, gold_geo["depth_points"])        # same float32 operations in the same order


def test_geometry_oracle_depth_gt_matches_reference_code(gold_geo):
    from oracle import geometry_oracle as geo
    W, H = int(gold_geo["cam"][0]), int(gold_geo["cam"][1])
    d = geo.compute_depth_gt(gold_geo["gt_points"], gold_geo["K"], gold_geo["gt_c2w"], H, W)
    assert d.shape == (H, W)
    assert U.rel_err(d, gold_geo["gt_depth"]) < 1e-5


@pytest.mark.parametrize("tag", ["lattice", "blob"])
def test_geometry_oracle_pca_matches_reference_code(gold_geo, tag):
    """align_principle_axes / transform_cameras (src/data/normalize.py) run by the fixture generator vs the oracle."""
    from oracle import geometry_oracle as geo
    T = geo.align_principle_axes(gold_geo[tag + "_points"])
    assert np.allclose(T, gold_geo[tag + "_pca_T"], atol=1e-6)
    med, cov = geo.median_cov(gold_geo[tag + "_points"])
    R = gold_geo[tag + "_pca_T"][:3, :3]
    assert np.allclose(-R @ med, gold_geo[tag + "_pca_T"][:3, 3], atol=1e-5)
    D = R @ cov @ R.T                                   # the reference's axes diagonalise the float64 covariance
    assert np.abs(D - np.diag(np.diag(D))).max() < 1e-5 * np.abs(D).max()
    assert D[0, 0] >= D[1, 1] >= D[2, 2] and np.linalg.det(R) > 0
