"""CPU tests of the oracle itself (it is unpinned by the reference — see oracle/gsplat_oracle.py header — so it
is pinned by self-consistency: finite differences, invariants, convergence)."""
import math

import numpy as np
import torch

from oracle import gsplat_oracle as orc
import gsb_testutil as U


def _setup(dtype=torch.float64):
    fp = U.frame()
    means, quats, scales, opac = U.map_tensors(fp, dtype)
    K = torch.tensor(fp.K, dtype=dtype)
    gt = torch.tensor(fp.src_depth, dtype=dtype)
    q0, t0 = U.offgrid_pose(fp, dtype)
    return fp, means, quats, scales, opac, K, gt, q0, t0


def test_pose_gradient_matches_finite_differences():
    fp, means, quats, scales, opac, K, gt, q0, t0 = _setup()
    q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
    tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q, t, means, quats, scales, opac, gt, K, fp.W, fp.H)
    tot.backward()

    def f(qv, tv):  # discrete skip/stop/tile decisions frozen at the centre point: the smooth piece
        with torch.no_grad():
            return float(orc.pose_loss(qv, tv, means, quats, scales, opac, gt, K, fp.W, fp.H, frozen_meta=meta)[0])
    eps = 1e-6
    for i in range(4):
        e = torch.zeros(4, dtype=torch.float64); e[i] = eps
        fd = (f(q0 + e, t0) - f(q0 - e, t0)) / (2 * eps)
        assert abs(fd - float(q.grad[i])) <= 1e-6 * max(1.0, abs(fd)), (i, fd, float(q.grad[i]))
    for i in range(3):
        e = torch.zeros(3, dtype=torch.float64); e[i] = eps
        fd = (f(q0, t0 + e) - f(q0, t0 - e)) / (2 * eps)
        assert abs(fd - float(t.grad[i])) <= 1e-6 * max(1.0, abs(fd)), (i, fd, float(t.grad[i]))


def test_normal_loss_gradient_matches_finite_differences():
    fp, means, quats, scales, opac, K, gt, q0, t0 = _setup()
    for dim in (1, -1):
        q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
        tot, *_, meta = orc.pose_loss(q, t, means, quats, scales, opac, gt, K, fp.W, fp.H, 0.7, 0.1, dim)
        tot.backward()
        eps = 1e-6
        e = torch.zeros(3, dtype=torch.float64); e[2] = eps
        with torch.no_grad():
            fp_ = float(orc.pose_loss(q0, t0 + e, means, quats, scales, opac, gt, K, fp.W, fp.H, 0.7, 0.1, dim, frozen_meta=meta)[0])
            fm_ = float(orc.pose_loss(q0, t0 - e, means, quats, scales, opac, gt, K, fp.W, fp.H, 0.7, 0.1, dim, frozen_meta=meta)[0])
        fd = (fp_ - fm_) / (2 * eps)
        assert abs(fd - float(t.grad[2])) <= 1e-5 * max(1.0, abs(fd))


def test_forward_invariants():
    fp, means, quats, scales, opac, K, gt, q0, t0 = _setup(torch.float32)
    with torch.no_grad():
        tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q0, t0, means, quats, scales, opac, gt, K, fp.W, fp.H)
    assert float(a.max()) <= 1.0 and float(a.min()) >= 0.0
    ids, fl, offs = meta["isect_ids"], meta["flatten_ids"], meta["isect_offsets"].reshape(-1)
    assert np.all(np.diff(ids) >= 0), "keys sorted"
    assert np.all(np.diff(offs) >= 0) and offs[0] == 0 and offs[-1] <= len(ids), "offsets monotone"
    assert int(meta["tiles_per_gauss"].sum()) == len(ids)
    # every entry of a tile's range carries that tile id
    tb = int(math.floor(math.log2(meta["tile_width"] * meta["tile_height"]))) + 1
    tile_of = (ids >> 32) & ((1 << tb) - 1)
    ends = np.concatenate([offs[1:], [len(ids)]])
    for t in (0, len(offs) // 2, len(offs) - 1):
        assert np.all(tile_of[offs[t]:ends[t]] == t)
    # depth > 0 wherever something was composited; zero-depth pixels have zero alpha
    dn, an = d.numpy(), a.numpy()
    assert np.all(dn[an > 0] > 0) and np.all(an[dn == 0] == 0)


def test_stable_ties_keep_emission_order():
    # two Gaussians with bit-identical depth in the same tile: the lower flat index must come first
    m2 = np.array([[[20.0, 20.0], [10.0, 10.0], [21.0, 19.0]]], dtype=np.float32)
    radii = np.array([[6, 6, 6]], dtype=np.int32)
    depths = np.array([[2.0, 1.0, 2.0]], dtype=np.float32)
    tpg, ids, fl = orc.isect_tiles(m2, radii, depths, 16, 4, 4)
    offs = orc.isect_offset_encode(ids, 1, 4, 4).reshape(-1)
    first_tile = fl[offs[0]:offs[1]]
    assert list(first_tile) == [1, 0, 2]


def test_isect_empty_and_offscreen():
    m2 = np.zeros((2, 5, 2), dtype=np.float32)
    radii = np.zeros((2, 5), dtype=np.int32)
    tpg, ids, fl = orc.isect_tiles(m2, radii, np.ones((2, 5), dtype=np.float32), 16, 3, 2)
    assert ids.size == 0 and fl.size == 0 and tpg.sum() == 0
    offs = orc.isect_offset_encode(ids, 2, 3, 2)
    assert offs.shape == (2, 2, 3) and np.all(offs == 0)


def test_pose_loop_converges_to_the_known_delta():
    torch.manual_seed(0)
    fp = U.frame(U.SMALL_CAM, stride=2, seed=7, trans_cm=1.0, rot_deg=0.5)
    dtype = torch.float32
    means, quats, scales, opac = U.map_tensors(fp, dtype)
    K = torch.tensor(fp.K, dtype=dtype)
    gt = U.oracle_gt_depth(fp, dtype)
    q0 = torch.tensor(orc.rotmat_to_quat(fp.tar_c2w[:3, :3]), dtype=dtype)
    t0 = torch.tensor(fp.tar_c2w[:3, 3], dtype=dtype)
    eT0, eR0 = orc.pose_errors(fp.tar_c2w, fp.src_c2w)
    res = orc.pose_opt_loop(q0, t0, means, quats, scales, opac, gt, K, fp.W, fp.H, c2w_gt=fp.src_c2w, max_steps=1000,
                            n_steps=60, early_stop=False)
    eT, eR = res["history"][-1][3], res["history"][-1][4]
    # the loss floor is set by the coarse splats of a 160x96 frame, so ask for a clear drop, not for zero
    assert res["history"][-1][0] < 0.9 * res["history"][0][0], "loss must drop"
    assert eT < 0.4 * eT0 and eR < 0.4 * eR0, (eT0, eT, eR0, eR)


def test_exposure_to_the_uncertain_constants():
    """the rasteriser's oracle restates gsplat 1.3.0 and cannot be pinned to
    the absent fork, so the two constants that neighbouring versions set differently are switches (oracle VARIANT, kernels
    -DGSB_RADIUS_FLOOR / -DGSB_CLAMP_ASYMMETRIC) and their effect on the synthetic frames is measured here:
      B2 radius floor 0.01 -> 0.1 (Inria): GsplatLoc's maps are isotropic, so b^2 - det sits under either floor for every
         Gaussian and v1 moves by sqrt(0.1) - sqrt(0.01) = 0.216 px^2: the 3-sigma radius of a few per cent of the Gaussians
         crosses an integer, their tile lists grow, pixels at the rim (alpha between 1/255 and exp(-4.5)) change;
      B1 asymmetric frustum clamp (gsplat >= 1.4.0): differs only for Gaussians whose centre projects more than 30 % outside
         the image -- none on these frames."""
    fp = U.frame(U.MID_CAM, stride=2, seed=3)
    dt = torch.float64
    means, quats, scales, opac = U.map_tensors(fp, dt)
    gt, K = torch.tensor(fp.src_depth, dtype=dt), torch.tensor(fp.K, dtype=dt)
    q0, t0 = U.offgrid_pose(fp, dt)

    def run():
        q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
        tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q, t, means, quats, scales, opac, gt, K, fp.W, fp.H)
        tot.backward()
        return dict(loss=float(tot), d=d.detach().numpy(), a=a.detach().numpy(), radii=meta["radii"].numpy(),
                    M=int(meta["flatten_ids"].size), gq=q.grad.numpy(), gt=t.grad.numpy())

    saved = dict(orc.VARIANT)
    try:
        base = run()
        orc.VARIANT["radius_floor"] = 0.1
        b2 = run()
        orc.VARIANT.update(saved)
        orc.VARIANT["clamp"] = "asymmetric"
        b1 = run()
    finally:
        orc.VARIANT.update(saved)
    vis = base["radii"] > 0
    moved = float((b2["radii"][vis] != base["radii"][vis]).mean())
    exposure = {"B2_gaussians_with_another_radius": moved, "B2_isects": (base["M"], b2["M"]),
                "B2_depth_rel": U.rel_err(b2["d"], base["d"]), "B2_alpha_rel": U.rel_err(b2["a"], base["a"]),
                "B2_loss_rel": abs(b2["loss"] - base["loss"]) / base["loss"],
                "B2_grad_q_rel": U.rel_err(b2["gq"], base["gq"]), "B2_grad_t_rel": U.rel_err(b2["gt"], base["gt"]),
                "B1_grad_rel": max(U.rel_err(b1["gq"], base["gq"]), U.rel_err(b1["gt"], base["gt"])),
                "B1_isects": (base["M"], b1["M"])}
    print(exposure)
    # B1: no Gaussian of these frames is clamped -> nothing moves at all
    assert b1["M"] == base["M"] and np.array_equal(b1["radii"], base["radii"]) and exposure["B1_grad_rel"] < 1e-12
    # B2: a minority of radii grow by one pixel; the render and the pose gradient move by far less than the parity
    # tolerances ask of the CUDA path (1e-4 / 1e-3) -- but not by nothing, which is why the constant is recorded
    assert 0.0 < moved < 0.5 and b2["M"] >= base["M"]
    assert np.all(b2["radii"][vis] - base["radii"][vis] >= 0) and np.all(b2["radii"][vis] - base["radii"][vis] <= 1)
    assert exposure["B2_depth_rel"] < 1e-3 and exposure["B2_alpha_rel"] < 1e-3
    assert exposure["B2_grad_q_rel"] < 2e-2 and exposure["B2_grad_t_rel"] < 2e-2
