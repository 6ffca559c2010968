"""Tile-striped single frame: the striping of the kernels and the two exact sums, with the ranks
EMULATED on one GPU (one striped tracker per rank on the same device, the all-reduce replaced by a sum over their
buffers) so that the test runs on a one-GPU box; and the same through torch.distributed / NCCL when two GPUs are there
(tools/run_stripe.sh runs that one under torchrun).  The bar: bit-identical to the unstriped tracker."""
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


def _make(gsb, fp, c2w0, **cfg):
    means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
    trk = gsb.PoseTracker(fp.W, fp.H, 1, gsb.TrackerConfig(**cfg))
    trk.set_map(means, quats, scales, opac)
    trk.set_observation(torch.as_tensor(fp.src_depth, dtype=torch.float32).cuda(), fp.K)
    trk.set_poses(c2w0, fp.src_c2w)
    return trk


@pytest.mark.parametrize("world", [2, 4])
def test_striped_ranks_reproduce_the_single_gpu_trajectory_bit_for_bit(gsb, world):
    from gsplatloc_b200.stripe import StripedTracker
    fp = U.frame(U.MID_CAM, stride=2, seed=13)
    q0, t0 = U.offgrid_pose(fp, torch.float64)
    c2w0 = orc.pose_to_c2w(q0, t0).numpy()
    n = 12
    ref = _make(gsb, fp, c2w0, early_stop=False)
    ref.run(n, poll_every=0, use_graph=False)
    ref_state, ref_hist = ref.states()[0], ref.history(0)
    ref_d, ref_a = ref.render()

    ranks = []

    def emulated_all_reduce(kind):
        def f(_):
            # (called once per rank and collective; the sum is formed when the last rank arrives -- every rank of the
            # emulation runs on the same stream order, so all inputs are complete by then)
            f.calls += 1
            if f.calls % world == 0:
                bufs = [getattr(r, kind) for r in ranks]
                total = torch.stack(bufs).sum(0)
                for b in bufs:
                    b.copy_(total)
        f.calls = 0
        return f

    ar_depth, ar_part = emulated_all_reduce("depth"), emulated_all_reduce("partials")
    stream = torch.cuda.Stream()
    for r in range(world):
        means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
        trk = gsb.PoseTracker(fp.W, fp.H, 1, gsb.TrackerConfig(early_stop=False), stream=stream)
        trk.set_map(means, quats, scales, opac)
        trk.set_observation(torch.as_tensor(fp.src_depth, dtype=torch.float32).cuda(), fp.K)
        trk.set_poses(c2w0, fp.src_c2w)
        st = StripedTracker(trk, r, world, all_reduce=lambda t: None)
        ranks.append(st)
    with torch.cuda.stream(stream):
        for _ in range(n):      # lock step over the emulated ranks: segment, "collective", segment, "collective", segment
            for st in ranks:
                st.segment(0)
            for st in ranks:
                ar_depth(st.depth)
            for st in ranks:
                st.segment(1)
            for st in ranks:
                ar_part(st.partials)
            for st in ranks:
                st.segment(2)
    for st in ranks:
        st.finish()
        s = st.trk.states()[0]
        assert np.array_equal(s["q"], ref_state["q"]) and np.array_equal(s["t"], ref_state["t"]), (st.rank, s["q"], ref_state["q"])
        assert np.float32(s["loss"]).view(np.uint32) == np.float32(ref_state["loss"]).view(np.uint32)
        assert np.array_equal(s["g_q"], ref_state["g_q"]) and np.array_equal(s["g_t"], ref_state["g_t"])
        assert np.array_equal(st.trk.history(0), ref_hist)
    # the summed image on every rank is the single-GPU render; each rank rendered about 1 / world of the intersections
    assert torch.equal(ranks[0].depth.view(fp.H, fp.W), ref_d[0])
    M = [st.trk.counters()["n_isects"] for st in ranks]
    assert sum(M) == ref.counters()["n_isects"] and max(M) < 1.5 * sum(M) / world
    for st in ranks:
        st.trk.close()
    ref.close()


def test_fused_exchange_reproduces_the_single_gpu_trajectory_bit_for_bit(gsb):
    """The fused exchange (kernels store their own tiles' pixels / partials into EVERY rank's buffers, no all-reduce): four
    emulated ranks on one device and one stream (stream order stands in for the barrier kernels, which need one GPU per
    rank: tools/stripe_parity.py --fused under torchrun)."""
    from gsplatloc_b200.stripe import StripedTracker
    world, n = 4, 10
    fp = U.frame(U.MID_CAM, stride=2, seed=17)
    q0, t0 = U.offgrid_pose(fp, torch.float64)
    c2w0 = orc.pose_to_c2w(q0, t0).numpy()
    ref = _make(gsb, fp, c2w0, early_stop=False)
    ref.run(n, poll_every=0, use_graph=False)
    ref_state, ref_hist = ref.states()[0], ref.history(0)
    stream = torch.cuda.Stream()
    ranks = []
    for r in range(world):
        means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
        trk = gsb.PoseTracker(fp.W, fp.H, 1, gsb.TrackerConfig(early_stop=False), stream=stream)
        trk.set_map(means, quats, scales, opac)
        trk.set_observation(torch.as_tensor(fp.src_depth, dtype=torch.float32).cuda(), fp.K)
        trk.set_poses(c2w0, fp.src_c2w)
        ranks.append(StripedTracker(trk, r, world, all_reduce=lambda t: None))
    for st in ranks:
        st.fuse_with_local_ranks(ranks)
        st.segment(-1)
    with torch.cuda.stream(stream):
        for _ in range(n):
            for k in (0, 1, 2):
                for st in ranks:
                    st.segment(k)
    for st in ranks:
        st.finish()
        s = st.trk.states()[0]
        assert np.array_equal(s["q"], ref_state["q"]) and np.array_equal(s["t"], ref_state["t"]), (st.rank, s["q"], ref_state["q"])
        assert np.array_equal(st.trk.history(0), ref_hist)
    d, _ = ref.render()
    assert torch.equal(ranks[2].depth.view(fp.H, fp.W), d[0])
    for st in ranks:
        st.trk.close()
    ref.close()


def test_stripe_argument_checks(gsb):
    from gsplatloc_b200.stripe import StripedTracker
    fp = U.frame(U.SMALL_CAM, stride=2, seed=1)
    trk = _make(gsb, fp, fp.tar_c2w, early_stop=False)
    with pytest.raises(ValueError):
        StripedTracker(trk, 0, 3)
    with pytest.raises(ValueError):
        StripedTracker(trk, 2, 2)
    one = StripedTracker(trk, 0, 1)      # world 1: the plain tracker, no collective
    one.run(3, poll_every=0)
    assert trk.states()[0]["iter"] == 3
    trk.close()


@pytest.mark.parametrize("fused", [False, True])
def test_two_hypotheses_and_the_early_stop_on_striped_ranks(gsb, fused):
    """C = 2 pose hypotheses of the striped frame with the early stop on (short patience so that it fires inside the run):
    every emulated rank must stop both cameras in the iteration the unstriped tracker stops them and hold the same states."""
    from gsplatloc_b200.stripe import StripedTracker
    world, n = 2, 40
    fp = U.frame(U.SMALL_CAM, stride=2, seed=23, trans_cm=1.0, rot_deg=0.5)
    p0 = fp.tar_c2w.astype(np.float64)
    p1 = p0.copy(); p1[:3, 3] += np.array([0.004, -0.003, 0.002])
    poses = np.stack([p0, p1])
    cfg = dict(early_stop=True, patience=4, warmup=3, max_steps=200, trans_lr=0.02, quat_lr=0.01)   # steps far too large: the loss bounces

    def make(stream=None):
        means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
        trk = gsb.PoseTracker(fp.W, fp.H, 2, gsb.TrackerConfig(**cfg), stream=stream)
        trk.set_map(means, quats, scales, opac)
        trk.set_observation(torch.as_tensor(fp.src_depth, dtype=torch.float32).cuda(), fp.K)
        trk.set_poses(poses, fp.src_c2w)
        return trk

    ref = make()
    ref.run(n, poll_every=0, use_graph=False)
    want = ref.states()
    assert any(s["stopped"] for s in want), "the early stop must fire inside the run for the test to mean something"
    stream = torch.cuda.Stream()
    ranks = [StripedTracker(make(stream), r, world, all_reduce=lambda t: None) for r in range(world)]
    if fused:
        for st in ranks:
            st.fuse_with_local_ranks(ranks)
    for st in ranks:
        st.segment(-1)

    def summed(kind):
        total = torch.stack([getattr(r, kind) for r in ranks]).sum(0)
        for r in ranks:
            getattr(r, kind).copy_(total)

    with torch.cuda.stream(stream):
        for _ in range(n):
            for st in ranks:
                st.segment(0)
            if not fused:
                summed("depth")
            for st in ranks:
                st.segment(1)
            if not fused:
                summed("partials")
            for st in ranks:
                st.segment(2)
    for st in ranks:
        st.finish()
        got = st.trk.states()
        for c in range(2):
            assert got[c]["stopped"] == want[c]["stopped"] and got[c]["iter"] == want[c]["iter"], (st.rank, c)
            assert np.array_equal(got[c]["q"], want[c]["q"]) and np.array_equal(got[c]["t"], want[c]["t"])
            assert got[c]["best_iter"] == want[c]["best_iter"] and got[c]["best_eT"] == want[c]["best_eT"]
        st.trk.close()
    ref.close()
