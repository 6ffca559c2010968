"""Host logic of the multi-GPU path on the CPU: world_size-2 ``gloo`` processes exercise the partitioning and the one
collective (the final gather of per-frame records) exactly as the NCCL ranks do on the GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsplatloc_b200 import sharding as sh


def test_frame_partition_is_exact_and_balanced():
    for n in (0, 1, 7, 500):
        for world in (1, 2, 4, 8):
            parts = [sh.frames_of_rank(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_hypothesis_partition_is_contiguous_and_exact():
    for n in (1, 3, 64, 65):
        for world in (1, 2, 4, 8):
            parts = [sh.hypotheses_of_rank(n, r, world) for r in range(world)]
            assert [i for p in parts for i in p] == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def _fake_state(unit):
    rng = np.random.default_rng(unit)
    return {"iter": 100 + unit, "stopped": unit % 2, "loss": float(rng.random()), "depth_loss": 0.1, "edge_loss": 0.2,
            "normal_loss": 0.0, "eT": 1e-3 * (unit + 1), "eR": 1e-2 * (unit + 1), "best_loss": float(rng.random()),
            "best_eT": 1e-3, "best_eR": 1e-2, "best_iter": 150, "q": rng.random(4), "t": rng.random(3)}


def _worker(rank, world, port, n_frames, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sh.frames_of_rank(n_frames, rank, world)
        rows = [sh.state_to_record(f, rank, _fake_state(f)) for f in mine]
        allrec = sh.gather_records(rows)
        np.save(os.path.join(out_dir, f"rec{rank}.npy"), allrec)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n_frames", [7, 1])
def test_gather_records_world2_gloo(tmp_path, n_frames):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_frames, str(tmp_path)), nprocs=world, join=True)
    recs = [np.load(tmp_path / f"rec{r}.npy") for r in range(world)]
    assert np.array_equal(recs[0], recs[1]), "every rank must hold the same gathered table"
    r = recs[0]
    assert r.shape == (n_frames, sh.RECORD_WIDTH)
    assert list(r[:, 0]) == list(range(n_frames)), "sorted by frame id"
    assert list(r[:, 1]) == [f % world for f in range(n_frames)], "round-robin ownership"
    for f in range(n_frames):
        assert np.allclose(r[f], sh.state_to_record(f, f % world, _fake_state(f)))
    err = sh.trajectory_errors(r)
    eT = np.array([1e-3 * (f + 1) for f in range(n_frames)])
    assert abs(err["ATE_rmse_m"] - np.sqrt((eT ** 2).mean())) < 1e-12
    best = sh.best_hypothesis(r, "loss")
    assert best == int(np.argmin([_fake_state(f)["loss"] for f in range(n_frames)]))


def test_gather_records_single_process_is_identity_sorted():
    rows = [sh.state_to_record(f, 0, _fake_state(f)) for f in (3, 1, 2)]
    r = sh.gather_records(rows)
    assert list(r[:, 0]) == [1, 2, 3]


def test_stripe_ownership_partitions_tiles_and_loss_blocks():
    """Row f3 host logic (no GPU): every tile has one owner, the two tiles of a horizontal pair share it, the four 32x8
    loss blocks inside a pair belong to the same rank, and the interleaving balances the ranks to within a few pairs --
    for the three benchmark shapes (75x43 tiles with an odd width, 40x30, 120x68) and world sizes 1 ... 8."""
    import numpy as np
    pytest = __import__("pytest")
    # the CUDA library is not needed for the partition functions, but the module imports the tracker: skip if it cannot
    stripe = pytest.importorskip("gsplatloc_b200.stripe")
    for W, H in ((1200, 680), (640, 480), (1920, 1080)):
        tw, th = (W + 15) // 16, (H + 15) // 16
        tx, ty = np.meshgrid(np.arange(tw), np.arange(th))
        for world in (1, 2, 4, 8):
            own = stripe.tile_owner(tx, ty, tw, world)
            assert own.min() >= 0 and own.max() == world - 1
            left = own[:, 0:tw - (tw % 2):2]
            right = own[:, 1:tw:2]
            assert np.array_equal(left[:, :right.shape[1]], right), "the tiles of a pair share their owner"
            share = np.bincount(own.ravel(), minlength=world) / own.size
            assert share.max() - share.min() < 0.02, (W, H, world, share)
            # loss blocks: 32x8 pixels; block (bx, by) lies in tiles (2bx, by // 2) and (2bx + 1, by // 2)
            gx, gy = (W + 31) // 32, (H + 7) // 8
            bx, by = np.meshgrid(np.arange(gx), np.arange(gy))
            lown = stripe.loss_block_owner(bx, by, tw, world)
            assert np.array_equal(lown, own[np.minimum(by // 2, th - 1), np.minimum(2 * bx, tw - 1)])
    with pytest.raises(ValueError):
        stripe.tile_owner(0, 0, 10, 3)
