"""GPU parity tests of the device-resident pose loop (the inner loop of Runner.train, gs_trainer_total.py:78-267)
against the CPU oracle: intermediates bit-exact, render 1e-4, pose gradient 1e-3, converged pose 1e-3 m."""
import math

import numpy as np
import pytest
import torch

import gsb_testutil as U
from oracle import gsplat_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gsb(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import gsplatloc_b200
    return gsplatloc_b200


def _tracker(gsb, fp, c2w0, gt, C=1, **cfg):
    means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
    trk = gsb.PoseTracker(fp.W, fp.H, C, gsb.TrackerConfig(**cfg))
    trk.set_map(means, quats, scales, opac)
    trk.set_observation(torch.as_tensor(gt, dtype=torch.float32).cuda(), fp.K)
    trk.set_poses(c2w0, fp.src_c2w)
    return trk


def _oracle_step(fp, c2w0, gt, dt=torch.float64):
    means, quats, scales, opac = U.map_tensors(fp, dt)
    q = torch.tensor(orc.rotmat_to_quat(c2w0[:3, :3]), dtype=dt).requires_grad_(True)
    t = torch.tensor(c2w0[:3, 3], dtype=dt).requires_grad_(True)
    out = orc.pose_loss(q, t, means, quats, scales, opac, torch.as_tensor(gt, dtype=dt), torch.tensor(fp.K, dtype=dt),
                        fp.W, fp.H)
    out[0].backward()
    return q, t, out


def _rendered_gt(gsb, fp):
    """observation = splat render of the source frame's own points from its own pose (dataset.py:362-372)"""
    means_s = torch.tensor(fp.src_points).cuda()
    Ns = means_s.shape[0]
    gtr = gsb.PoseTracker(fp.W, fp.H, 1)
    gtr.set_map(means_s, torch.tensor([[1.0, 0, 0, 0]]).repeat(Ns, 1).cuda(), torch.tensor(fp.src_scales).cuda(),
                torch.ones(Ns).cuda())
    gtr.set_observation(torch.zeros(fp.H, fp.W).cuda(), fp.K)
    gtr.set_poses(fp.src_c2w)
    gtr.forward()
    gt, _ = gtr.render()
    gtr.close()
    return gt[0].clone()


def _start_pose(fp):
    q0, t0 = U.offgrid_pose(fp, torch.float64)
    return orc.pose_to_c2w(q0, t0).numpy()


@pytest.mark.parametrize("cam,stride,seed", [(U.SMALL_CAM, 2, 42), (U.MID_CAM, 2, 3), (U.SMALL_CAM, 1, 5)])
def test_one_iteration_matches_oracle(gsb, cam, stride, seed):
    fp = U.frame(cam, stride=stride, seed=seed)
    c2w0 = _start_pose(fp)
    trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
    trk.run(1, poll_every=0, use_graph=False)
    st = trk.states()[0]
    q, t, (tot, Ld, Ls, Ln, d, a, meta) = _oracle_step(fp, c2w0, fp.src_depth)
    d_gpu, a_gpu = trk.render()
    assert U.rel_err(d_gpu[0].cpu().numpy(), d.detach().numpy()) < 1e-4
    assert U.rel_err(a_gpu[0].cpu().numpy(), a.detach().numpy()) < 1e-4
    assert abs(st["depth_loss"] - float(Ld)) / float(Ld) < 1e-4
    assert abs(st["edge_loss"] - float(Ls)) / float(Ls) < 1e-4
    assert abs(st["loss"] - float(tot)) / float(tot) < 1e-4
    assert U.rel_err(st["g_q"], q.grad.numpy()) < 1e-3, (st["g_q"], q.grad)
    assert U.rel_err(st["g_t"], t.grad.numpy()) < 1e-3, (st["g_t"], t.grad)
    eT, eR = orc.pose_errors(c2w0, fp.src_c2w)
    assert abs(st["eT"] - eT) < 1e-5 and abs(st["eR"] - eR) < 1e-2
    trk.close()


def test_intermediates_bit_exact_given_the_kernel_projection(gsb):
    """tile ids / sorted order / offsets: bit-exact against the oracle's intersection + stable sort when both
    start from the SAME projected means2d / radii / depths (the kernel's own records)."""
    fp = U.frame(U.MID_CAM, stride=2, seed=9)
    c2w0 = _start_pose(fp)
    trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
    trk.forward()
    fl, off, rec = trk.debug_tensors()
    rec = rec.cpu().numpy()
    radii = rec[:, 7].view(np.int32).copy()[None]
    m2 = rec[None, :, 0:2].copy()
    m2[radii == 0] = 0
    dep = rec[None, :, 2].copy()
    tw, th = math.ceil(fp.W / 16), math.ceil(fp.H / 16)
    tpg_o, ids_o, fl_o = orc.isect_tiles(m2, radii, dep, 16, tw, th)
    off_o = orc.isect_offset_encode(ids_o, 1, tw, th).reshape(-1)
    assert trk.counters()["n_isects"] == ids_o.size
    assert np.array_equal(fl.cpu().numpy(), fl_o)
    assert np.array_equal(off.cpu().numpy()[:-1], off_o) and int(off[-1]) == ids_o.size
    # and the projection itself agrees with the oracle's within fp32 rounding
    means, quats, scales, opac = U.map_tensors(fp, torch.float64)
    view = torch.linalg.inv(torch.tensor(c2w0))[None]
    r_o, m_o, d_o, c_o = orc.project(means, quats, scales, view, torch.tensor(fp.K, dtype=torch.float64)[None], fp.W, fp.H)
    both = (radii[0] > 0) & (r_o[0].numpy() > 0)
    assert both.mean() > 0.5 and ((radii[0] > 0) != (r_o[0].numpy() > 0)).mean() < 1e-3
    assert np.abs(radii[0][both] - r_o[0].numpy()[both]).max() <= 1
    np.testing.assert_allclose(rec[both, 0:2], m_o[0].numpy()[both], rtol=1e-5, atol=1e-3)
    np.testing.assert_allclose(rec[both, 2], d_o[0].numpy()[both], rtol=1e-5)
    np.testing.assert_allclose(rec[both, 4:7], c_o[0].numpy()[both], rtol=2e-4, atol=1e-7)
    trk.close()


def test_loop_tracks_the_oracle_loop(gsb):
    """20 Adam iterations: the pose trajectory stays on the oracle's (fp32 oracle, same arithmetic order not
    guaranteed, so the bound is loose on the pose and tight on the loss)."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=7, trans_cm=1.0, rot_deg=0.5)
    gt = U.oracle_gt_depth(fp, torch.float32).numpy()
    c2w0 = fp.tar_c2w.astype(np.float64)
    n = 20
    trk = _tracker(gsb, fp, c2w0, gt, early_stop=False, max_steps=1000)
    trk.run(n, poll_every=0, use_graph=True)
    hist = trk.history(0)
    st = trk.states()[0]
    dt = torch.float32
    means, quats, scales, opac = U.map_tensors(fp, dt)
    res = orc.pose_opt_loop(torch.tensor(orc.rotmat_to_quat(c2w0[:3, :3]), dtype=dt), torch.tensor(c2w0[:3, 3], dtype=dt),
                            means, quats, scales, opac, torch.tensor(gt), torch.tensor(fp.K), fp.W, fp.H,
                            c2w_gt=fp.src_c2w, max_steps=1000, n_steps=n, early_stop=False)
    assert hist.shape[0] == n
    lo = np.array([h[0] for h in res["history"]])
    assert np.abs(hist[:, 0] - lo).max() / lo.max() < 2e-3, (hist[:, 0], lo)
    assert np.abs(st["t"] - res["t"].numpy()).max() < 2e-4
    assert np.abs(st["q"] - res["quat"].numpy()).max() < 2e-4
    trk.close()


def test_graph_and_eager_are_identical_and_deterministic(gsb):
    fp = U.frame(U.SMALL_CAM, stride=2, seed=21)
    c2w0 = _start_pose(fp)
    outs = []
    for use_graph in (False, True, True):
        trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
        trk.run(15, poll_every=5, use_graph=use_graph)
        st = trk.states()[0]
        outs.append(np.concatenate([st["q"], st["t"], [st["loss"]]]))
        trk.close()
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
    assert np.array_equal(outs[1].view(np.uint32), outs[2].view(np.uint32))


def test_unrolled_chained_graph_is_bit_identical_to_plain_launches(gsb, monkeypatch):
    """gsb_tracker_run replays ten iterations per graph launch between two polls, the kernels chained by programmatic
    dependent launch (kernels.h).  27 iterations = two ten-iteration graphs + seven single ones; against eager launches,
    against GSB_PDL=0 (plain stream order) and against GSB_GRAPH_UNROLL=1 (one iteration per graph): the same bits, the
    same iteration counts.  The frame is small on purpose: every kernel of an iteration fits the GPU at once, the case in
    which a wrongly ordered chain would let a block read what its grand-predecessor is still writing."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=21)
    c2w0 = _start_pose(fp)

    def run(use_graph, env):
        for k in ("GSB_PDL", "GSB_GRAPH_UNROLL"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)     # (the switches are read at creation)
        out = trk.run(27, poll_every=0, use_graph=use_graph)
        st = trk.states()[0]
        trk.close()
        assert out["iters"] == 27 and st["iter"] == 27
        return np.concatenate([st["q"], st["t"], st["g_q"], st["g_t"], [st["loss"]]]).astype(np.float32).view(np.uint32)

    ref = run(False, {"GSB_PDL": "0"})
    for use_graph, env in ((True, {}), (False, {}), (True, {"GSB_PDL": "0"}), (True, {"GSB_GRAPH_UNROLL": "1"}),
                           (True, {"GSB_GRAPH_UNROLL": "4"})):
        assert np.array_equal(run(use_graph, env), ref), (use_graph, env)


def test_early_stop_inside_an_unrolled_graph_stops_at_the_same_iteration(gsb, monkeypatch):
    """The stop flag is raised on the device in the middle of a ten-iteration graph: the remaining iterations of the graph
    are no-ops, and the pose, the best pose and the iteration counter equal those of one-iteration graphs polled every
    iteration."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=17, trans_cm=0.5, rot_deg=0.2)
    outs = []
    for env, poll in (({"GSB_GRAPH_UNROLL": "1"}, 1), ({}, 0)):
        monkeypatch.delenv("GSB_GRAPH_UNROLL", raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        # (large learning rates make the loss oscillate so that the patience rule fires early)
        trk = _tracker(gsb, fp, fp.tar_c2w, fp.src_depth, max_steps=300, early_stop=True, patience=12, warmup=10,
                       trans_lr=2e-2, quat_lr=1e-2)
        trk.run(300, poll_every=poll, use_graph=True)
        st = trk.states()[0]
        trk.close()
        outs.append((st["stopped"], st["iter"], st["best_iter"],
                     np.concatenate([st["q"], st["t"], st["best_q"], st["best_t"]]).astype(np.float32).view(np.uint32)))
    assert outs[0][0] == 1 and outs[0][1] < 300, outs[0][:3]         # it did stop early
    assert outs[0][:3] == outs[1][:3]
    assert np.array_equal(outs[0][3], outs[1][3])


def test_converges_to_ground_truth_pose(gsb):
    """Converged ATE within 1e-3 m of the known synthetic delta (BASELINE.json north_star)."""
    fp = U.frame(U.MID_CAM, stride=1, seed=31, trans_cm=1.5, rot_deg=0.7)
    gt = _rendered_gt(gsb, fp)
    trk = _tracker(gsb, fp, fp.tar_c2w, gt, max_steps=1000, early_stop=True, patience=100)
    eT0, eR0 = orc.pose_errors(fp.tar_c2w, fp.src_c2w)
    res = trk.run(1000, poll_every=50)
    st = trk.states()[0]
    assert st["best_eT"] < 1e-3 and st["best_eT"] < 0.1 * eT0, (eT0, st["best_eT"])
    assert st["best_eR"] < 0.05 and st["best_eR"] < 0.2 * eR0, (eR0, st["best_eR"])
    assert st["stopped"] == 1 and res["iters"] < 1000, "the early stop (gs_trainer_total.py:160-185) must fire"
    trk.close()


def test_early_stop_bookkeeping_matches_reference_rule(gsb):
    """step > warmup: best tracking; counter >= patience: stop BEFORE the optimiser step (gs_trainer_total.py:160-185,
    245-263). Replays the rule on the recorded loss history."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=17, trans_cm=0.5, rot_deg=0.2)
    # large learning rates make the loss oscillate so that the patience rule fires early
    trk = _tracker(gsb, fp, fp.tar_c2w, fp.src_depth, max_steps=300, early_stop=True, patience=12, warmup=10,
                   trans_lr=2e-2, quat_lr=1e-2)
    trk.run(300, poll_every=10)
    st = trk.states()[0]
    hist = trk.history(0)
    best, counter, stop_at, best_iter = np.inf, 0, None, -1
    for step, row in enumerate(hist):
        if step > 10:
            if row[0] < best:
                best, counter, best_iter = row[0], 0, step
            else:
                counter += 1
        if counter >= 12:
            stop_at = step
            break
    assert stop_at is not None and st["stopped"] == 1
    assert hist.shape[0] == stop_at + 1
    assert st["best_iter"] == best_iter and st["best_loss"] == np.float32(best)
    assert st["adam_step"] == stop_at, "the gradient of the stopping iteration is discarded"
    trk.close()


def test_hypotheses_are_independent(gsb):
    """C pose hypotheses through the camera axis give exactly what C single-camera runs give."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=23)
    rng = np.random.default_rng(0)
    poses = []
    for c in range(3):
        p = _start_pose(fp).copy()
        p[:3, 3] += rng.normal(0, 0.01, 3)
        poses.append(p)
    multi = _tracker(gsb, fp, np.stack(poses), fp.src_depth, C=3, early_stop=False)
    multi.run(8, poll_every=0)
    ms = multi.states()
    for c in range(3):
        one = _tracker(gsb, fp, poses[c], fp.src_depth, early_stop=False)
        one.run(8, poll_every=0)
        s = one.states()[0]
        assert np.allclose(ms[c]["q"], s["q"], atol=1e-6) and np.allclose(ms[c]["t"], s["t"], atol=1e-6)
        assert abs(ms[c]["loss"] - s["loss"]) < 1e-6 * max(1.0, s["loss"])
        one.close()
    multi.close()


def test_full_size_properties_replica_200k(gsb):
    """BASELINE.json config 2 (1200x680, ~200k Gaussians): too big for the oracle, so check size-independent
    properties: sorted tile ranges, offsets monotone and closed, alpha in [0,1], depth within the scene range,
    loss decreasing, repeatable bits."""
    from gsplatloc_b200 import synthetic as syn
    fp = syn.make_frame_pair(syn.REPLICA, stride=2, seed=42)
    assert 190_000 < fp.tar_points.shape[0] < 210_000
    losses = []
    gt = _rendered_gt(gsb, fp)
    for rep in range(2):
        trk = _tracker(gsb, fp, fp.tar_c2w, gt, early_stop=False)
        trk.run(100, poll_every=0)
        st = trk.states()[0]
        losses.append(st["loss"])
        if rep == 0:
            hist = trk.history(0)
            fl, off, rec = trk.debug_tensors()
            M = trk.counters()["n_isects"]
            off = off.cpu().numpy().astype(np.int64)
            assert off[0] == 0 and off[-1] == M and np.all(np.diff(off) >= 0)
            dbits = rec[:, 2].contiguous().view(torch.int32)[fl.long()].cpu().numpy().astype(np.int64)
            tile_of = np.repeat(np.arange(off.size - 1), np.diff(off))
            key = (tile_of << 32) | dbits
            assert np.all(np.diff(key) >= 0), "every tile range sorted by depth bits"
            d, a = trk.render()
            assert float(a.min()) >= 0 and float(a.max()) <= 1.0
            dd = d[a > 0.5]
            assert float(dd.min()) > 0.2 and float(dd.max()) < 12.0
            assert hist[-1, 0] < hist[0, 0] and hist[-1, 4] < hist[0, 4], "loss and eT go down"
        trk.close()
    assert np.float32(losses[0]).view(np.uint32) == np.float32(losses[1]).view(np.uint32)


@pytest.mark.parametrize("mode,dim", [("pixel", -1), ("reference", 1)])
def test_one_iteration_with_normal_term_matches_oracle(gsb, mode, dim):
    """BASELINE.json configs[3]: depth + edge + normal loss (lambda = 0.7 / 0.2 / 0.1), pose gradient within 1e-3."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=11)
    c2w0 = _start_pose(fp)
    trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False, depth_lambda=0.7, normal_lambda=0.1, normal_mode=mode)
    trk.run(1, poll_every=0, use_graph=False)
    st = trk.states()[0]
    dt = torch.float64
    means, quats, scales, opac = U.map_tensors(fp, dt)
    q = torch.tensor(orc.rotmat_to_quat(c2w0[:3, :3]), dtype=dt).requires_grad_(True)
    t = torch.tensor(c2w0[:3, 3], dtype=dt).requires_grad_(True)
    tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q, t, means, quats, scales, opac, torch.as_tensor(fp.src_depth, dtype=dt),
                                                 torch.tensor(fp.K, dtype=dt), fp.W, fp.H, 0.7, 0.1, dim)
    tot.backward()
    assert abs(st["normal_loss"] - float(Ln)) / abs(float(Ln)) < 1e-4, (st["normal_loss"], float(Ln))
    assert abs(st["loss"] - float(tot)) / float(tot) < 1e-4
    assert U.rel_err(st["g_q"], q.grad.numpy()) < 1e-3, (st["g_q"], q.grad)
    assert U.rel_err(st["g_t"], t.grad.numpy()) < 1e-3, (st["g_t"], t.grad)
    assert trk.counters()["launches_per_iter"] == (7 if mode == "reference" else 6)
    trk.close()


@pytest.mark.parametrize("tag,dl,nl", [("A", 0.8, 0.0), ("B", 0.7, 0.1)])
def test_tracker_follows_the_reference_loop_fixture(gsb, tag, dl, nl):
    """tests/golden/ref_loop_small.npz: GsplatLoc's own loop code (pose module, losses, Adam, schedulers, metrics)
    run step by step on the CPU (make_golden.py).  The CUDA tracker must walk the same trajectory: loss within 1e-4
    relative, pose gradients within 1e-3 relative, parameters within 5e-6 absolute after every Adam step."""
    import os
    gold = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_loop_small.npz")))
    W, H = int(gold["cam"][0]), int(gold["cam"][1])
    N = gold["means"].shape[0]
    trk = gsb.PoseTracker(W, H, 1, gsb.TrackerConfig(max_steps=int(gold["max_steps"]), early_stop=False,
                                                     depth_lambda=dl, normal_lambda=nl, normal_mode="reference"))
    trk.set_map(torch.tensor(gold["means"]).cuda(), torch.tensor([[1.0, 0, 0, 0]]).repeat(N, 1).cuda(),
                torch.tensor(gold["scales"]).cuda(), torch.ones(N).cuda())
    trk.set_observation(torch.tensor(gold["gt_depth"]).cuda(), gold["K"])
    trk.set_poses(gold["c2w_init"], gold["c2w_gt"])
    n = gold[tag + "_quat"].shape[0]
    for s in range(n):
        st0 = trk.states()[0]
        assert np.allclose(st0["q"], gold[tag + "_quat"][s], atol=5e-6), (s, st0["q"], gold[tag + "_quat"][s])
        assert np.allclose(st0["t"], gold[tag + "_t"][s], atol=5e-6)
        trk.run(1, poll_every=0, use_graph=(s % 2 == 1))
        st = trk.states()[0]
        assert abs(st["loss"] - gold[tag + "_total"][s]) / gold[tag + "_total"][s] < 1e-4
        assert abs(st["depth_loss"] - gold[tag + "_Ld"][s]) / gold[tag + "_Ld"][s] < 1e-4
        assert abs(st["edge_loss"] - gold[tag + "_Ls"][s]) / gold[tag + "_Ls"][s] < 1e-4
        if nl:
            assert abs(st["normal_loss"] - gold[tag + "_Ln"][s]) / gold[tag + "_Ln"][s] < 1e-4
        assert U.rel_err(st["g_q"], gold[tag + "_g_quat"][s]) < 1e-3
        assert U.rel_err(st["g_t"], gold[tag + "_g_t"][s]) < 1e-3
        assert abs(st["eT"] - gold[tag + "_eT"][s]) < 1e-5 and abs(st["eR"] - gold[tag + "_eR"][s]) < 1e-2
        assert abs(st["lr_q"] - gold[tag + "_lr_q"][min(s + 1, n - 1)]) < 1e-9 or s == n - 1
        if s == 0:
            d, a = trk.render()
            assert U.rel_err(d[0].cpu().numpy(), gold[tag + "_first_depth"]) < 1e-4
            assert U.rel_err(a[0].cpu().numpy(), gold[tag + "_first_alpha"]) < 1e-4
    stf = trk.states()[0]
    assert np.allclose(stf["q"], gold[tag + "_final_quat"], atol=5e-6)
    assert np.allclose(stf["t"], gold[tag + "_final_t"], atol=5e-6)
    trk.close()


def test_one_iteration_anisotropic_map_matches_oracle(gsb):
    """General Gaussians (random unit quaternions, three different scales): the tracker's lift takes its general path
    (the isotropic short cut of raster_bwd.cu::lift_to_pose is off); loss within 1e-4, pose gradient within 1e-3."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=21)
    c2w0 = _start_pose(fp)
    g = torch.Generator().manual_seed(5)
    N = fp.tar_points.shape[0]
    quats = torch.nn.functional.normalize(torch.randn(N, 4, generator=g, dtype=torch.float64), dim=-1)
    scales = torch.tensor(fp.scales, dtype=torch.float64) * (0.6 + 0.8 * torch.rand(N, 3, generator=g, dtype=torch.float64))
    means = torch.tensor(fp.tar_points, dtype=torch.float64)
    opac = 0.3 + 0.7 * torch.rand(N, generator=g, dtype=torch.float64)
    trk = gsb.PoseTracker(fp.W, fp.H, 1, gsb.TrackerConfig(early_stop=False))
    trk.set_map(means.float().cuda(), quats.float().cuda(), scales.float().cuda(), opac.float().cuda())
    trk.set_observation(torch.tensor(fp.src_depth).cuda(), fp.K)
    trk.set_poses(c2w0, fp.src_c2w)
    trk.run(1, poll_every=0, use_graph=False)
    st = trk.states()[0]
    dt = torch.float64
    q = torch.tensor(orc.rotmat_to_quat(c2w0[:3, :3]), dtype=dt).requires_grad_(True)
    t = torch.tensor(c2w0[:3, 3], dtype=dt).requires_grad_(True)
    # the oracle sees the fp32-rounded map the kernel sees
    tot, Ld, Ls, Ln, d, a, meta = orc.pose_loss(q, t, means.float().to(dt), quats.float().to(dt), scales.float().to(dt),
                                                 opac.float().to(dt), torch.tensor(fp.src_depth, dtype=dt),
                                                 torch.tensor(fp.K, dtype=dt), fp.W, fp.H)
    tot.backward()
    d_gpu, a_gpu = trk.render()
    assert U.rel_err(d_gpu[0].cpu().numpy(), d.detach().numpy()) < 1e-4
    assert abs(st["loss"] - float(tot)) / float(tot) < 1e-4
    assert U.rel_err(st["g_q"], q.grad.numpy()) < 1e-3, (st["g_q"], q.grad)
    assert U.rel_err(st["g_t"], t.grad.numpy()) < 1e-3, (st["g_t"], t.grad)
    trk.close()


def test_bin_overflow_is_caught_and_the_run_repeated(gsb):
    """A tile bin that fills up must not corrupt anything silently: gsb_tracker_run notices the flag, restores the poses
    of the start of the run, doubles the bins and repeats -- the result is exactly that of an undisturbed tracker."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=42)
    c2w0 = _start_pose(fp)
    ref = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
    ref.run(30, poll_every=0)
    want = ref.states()[0]
    ref.close()
    for poll in (0, 10):                      # flag seen at the end of the run / at a poll point
        trk = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
        trk.debug_limit_bins(8)
        cap0 = trk.counters()["capacity"]
        trk.run(30, poll_every=poll)
        got = trk.states()[0]
        assert trk.counters()["capacity"] > cap0
        assert got["iter"] == want["iter"] == 30
        assert np.array_equal(np.asarray(got["q"]), np.asarray(want["q"]))
        assert np.array_equal(np.asarray(got["t"]), np.asarray(want["t"]))
        assert got["loss"] == want["loss"]
        trk.close()


@pytest.mark.parametrize("chain", ["last", "cv"])
def test_chained_walk_tracks_three_frames(gsb, chain):
    """Row f4 end to end on the device: ``sharding.track_frames(chain=...)`` places the map of every loop at the TRACKED
    pose of the loop before (model.py:41-77 bookkeeping on the host, everything else on the GPU); after three chained
    loops of 1.5 cm / 0.7 deg steps the accumulated drift stays under 3 mm."""
    from gsplatloc_b200 import geometry as geo, sharding, synthetic as syn
    cam, s = U.MID_CAM, 1
    W, H = cam["W"], cam["H"]
    K = np.array([[cam["fx"], 0, cam["cx"]], [0, cam["fy"], cam["cy"]], [0, 0, 1]], dtype=np.float64)
    rng = np.random.default_rng(31)
    scene, pose = syn.make_scene(rng), syn.make_pose(rng)
    poses = [pose]
    for i in range(3):
        poses.append(syn.perturb_pose(poses[-1], np.random.default_rng(100 + i), 1.5, 0.7))
    depths = [torch.tensor(syn.raycast_depth(scene, p, K, W, H)).cuda() for p in poses]
    Kt, K32 = torch.tensor(K.astype(np.float32)), K.astype(np.float32)
    eye, Ks = torch.eye(4)[None].cuda(), Kt[None].cuda()

    def load_frame(f, c2w_map):
        c2w_map = poses[f].astype(np.float32) if c2w_map is None else c2w_map
        pts = geo.depth_to_points(depths[f], Kt, c2w=torch.tensor(c2w_map)).contiguous()
        n = pts.shape[0]
        quats = torch.zeros((n, 4), device="cuda"); quats[:, 0] = 1.0
        cam_pts = geo.depth_to_points(depths[f + 1], Kt).contiguous()
        obs = geo.compute_depth_gt(cam_pts, None, Ks, eye, H, W, reference_faithful=False).contiguous()
        return dict(means=pts, quats=quats, scales=geo.init_gs_scales(pts, 5, reference_faithful=False),
                    opacities=torch.ones(n, device="cuda"), depth=obs, K=K32,
                    c2w_init=poses[f].astype(np.float32), c2w_gt=poses[f + 1].astype(np.float32))

    trk = gsb.PoseTracker(W, H, 1, gsb.TrackerConfig(max_steps=1000, early_stop=True, patience=100))
    res = sharding.track_frames([0, 1, 2], load_frame, trk, chain=chain)
    eT = res.records[:, sharding.RECORD_FIELDS.index("eT")]
    assert res.records.shape[0] == 3 and res.iters > 300
    assert eT.max() < 3e-3, eT                         # error against the ABSOLUTE ground truth: accumulated drift
    trk.close()


def test_bins_sized_for_another_pose_are_regrown_without_a_counting_pass(gsb):
    """Once the bins are sized for a map size, new poses (and the next frame of a sequence) skip the counting pass and its
    host synchronisation; bins that then prove too small are caught by the overflow flag -- ``forward`` recounts and
    renders again, ``run`` restores / re-sizes / repeats -- and the results are those of a freshly sized tracker."""
    fp = U.frame(U.SMALL_CAM, stride=2, seed=42)
    c2w0 = _start_pose(fp)
    far = c2w0.copy()
    far[:3, 3] -= 6.0 * far[:3, 2]            # six metres back along the optical axis: every Gaussian lands in a few tiles... or none
    fresh = _tracker(gsb, fp, c2w0, fp.src_depth, early_stop=False)
    fresh.forward()
    d_ref, a_ref = fresh.render()
    fresh.run(12, poll_every=0)
    want = fresh.states()[0]
    fresh.close()

    trk = _tracker(gsb, fp, far, fp.src_depth, early_stop=False)
    trk.forward()                             # sizes the bins at the far pose
    trk.set_poses(c2w0, fp.src_c2w)           # no counting pass now
    trk.forward()
    d, a = trk.render()
    assert torch.equal(d, d_ref) and torch.equal(a, a_ref)
    trk.set_poses(far, fp.src_c2w); trk.forward()
    trk.debug_limit_bins(8)                   # and the loop: bins far too small for the near pose
    trk.set_poses(c2w0, fp.src_c2w)
    trk.run(12, poll_every=0)
    got = trk.states()[0]
    assert np.array_equal(got["q"], want["q"]) and np.array_equal(got["t"], want["t"]) and got["loss"] == want["loss"]
    tiles = math.ceil(fp.W / 16) * math.ceil(fp.H / 16)
    assert trk.counters()["capacity"] > 8 * tiles      # the limited bins were re-sized by the repeat
    trk.close()
