#!/usr/bin/env python
"""Generates ``tests/golden/ref_loop_small.npz`` by running GsplatLoc's OWN Python code for the hot loop
(``/root/reference/src/my_gsplat/{model,loss,geometry,transform}.py``, ``src/eval/utils.py``) on a small seeded
synthetic frame, in the build container (CPU).  Run it from the repo root:

    python tests/golden/make_golden.py

What the fixture pins (reference code executed unmodified, see ``ref_shims.py`` for the import shims):
  * ``CameraOptModule_quat_tans``: quaternion initialisation, forward() -> c2w, the two Adam optimisers with
    coupled weight decay (model.py:27-116) and the ExponentialLR schedulers (gs_trainer_total.py:65-72,265-267);
  * ``GSModel.forward`` (model.py:180-215): the exact keyword set handed to ``rasterization`` (sigmoid of
    logit(1) opacities, SH-degree-1 colours, ``torch.linalg.inv(c2w)``), RGB+ED slicing (gs_trainer_total.py:104-105);
  * the trainer's mask and the three losses (gs_trainer_total.py:109-150; loss.py:10-101; geometry.py:138-197);
  * ``calculate_translation_error`` / ``calculate_rotation_error`` (eval/utils.py:122-168).
What it cannot pin: ``gsplat.rasterization`` itself (absent submodule) is the CPU oracle here, and the five kornia
functions are restated (ref_shims.py).  So the rendered depth inside the fixture is the oracle's; the fixture is a
regression pin for the rasteriser and a reference pin for everything around it.

The loop body below transcribes the call sequence of ``Runner.train`` (gs_trainer_total.py:78-152,263-267) — the
Runner class itself needs wandb, a dataset on disk and a viewer, none of which exist here.
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


def run_loop(ref, fp, c2w0, n_steps, max_steps, depth_lambda, normal_lambda):
    model, loss, geom = ref["model"], ref["loss"], ref["geometry"]
    tar_points = torch.tensor(fp.tar_points, dtype=torch.float32)
    colors = torch.full((tar_points.shape[0], 3), 0.5)
    gs_splats = model.GSModel(tar_points, colors)
    # scales are an INPUT of the hot path; use the synthetic generator's (true RMS neighbour distance) instead of
    # init_gs_scales' sqrt(mean(d^4)) so that the GPU tests can feed the same numbers
    gs_splats.scales = torch.tensor(fp.scales, dtype=torch.float32)
    depths_gt = torch.tensor(fp.src_depth, dtype=torch.float32)[None, :, :, None]       # [1,H,W,1]
    Ks = torch.tensor(fp.K, dtype=torch.float32)[None]
    camera_opt = model.CameraOptModule_quat_tans(torch.tensor(c2w0, dtype=torch.float32))
    schedulers = [torch.optim.lr_scheduler.ExponentialLR(camera_opt.optimizers[0], gamma=0.2 ** (1.0 / max_steps)),
                  torch.optim.lr_scheduler.ExponentialLR(camera_opt.optimizers[1], gamma=0.2 ** (1.0 / max_steps))]
    rows = {k: [] for k in ("quat", "t", "c2w", "total", "Ld", "Ls", "Ln", "g_quat", "g_t", "eT", "eR", "lr_q", "lr_t")}
    first = {}
    gt_pose = torch.tensor(fp.src_c2w, dtype=torch.float32)
    for step in range(n_steps):
        camera_opt.optimizer_clean()
        cur_c2w = camera_opt()
        renders, alphas, info = gs_splats(camtoworlds=cur_c2w.unsqueeze(0), Ks=Ks, width=fp.W, height=fp.H)
        assert renders.shape[-1] == 4
        depths = renders[..., 3:4]
        depths.retain_grad()
        non_zero_depth_mask = (depths != 0).float()
        depth_loss = loss.compute_depth_loss(depths * non_zero_depth_mask, depths_gt * non_zero_depth_mask, loss_type="l1")
        silhouette_loss = loss.compute_silhouette_loss(depths * non_zero_depth_mask, depths_gt * non_zero_depth_mask,
                                                       loss_type="l1")
        total_loss = depth_loss * depth_lambda + silhouette_loss * (1 - depth_lambda - normal_lambda)
        normal_loss = torch.zeros(())
        if normal_lambda != 0.0:
            normal_loss = loss.compute_normal_consistency_loss((depths_gt * non_zero_depth_mask)[0, :, :, 0],
                                                               (depths * non_zero_depth_mask)[0, :, :, 0],
                                                               K=Ks.squeeze(0), loss_type="cosine")
            total_loss = total_loss + normal_loss * normal_lambda
        total_loss.backward(retain_graph=True)
        with torch.no_grad():
            eT = ref["eval_utils"].calculate_translation_error(cur_c2w, gt_pose)
            eR = ref["eval_utils"].calculate_rotation_error(cur_c2w, gt_pose)
        rows["quat"].append(camera_opt.quat_cur.detach().numpy().copy())
        rows["t"].append(camera_opt.t_cur.detach().numpy().copy())
        rows["c2w"].append(cur_c2w.detach().numpy().copy())
        rows["total"].append(float(total_loss)); rows["Ld"].append(float(depth_loss))
        rows["Ls"].append(float(silhouette_loss)); rows["Ln"].append(float(normal_loss))
        rows["g_quat"].append(camera_opt.quat_cur.grad.numpy().copy())
        rows["g_t"].append(camera_opt.t_cur.grad.numpy().copy())
        rows["eT"].append(eT); rows["eR"].append(eR)
        rows["lr_q"].append(camera_opt.optimizers[0].param_groups[0]["lr"])
        rows["lr_t"].append(camera_opt.optimizers[1].param_groups[0]["lr"])
        if step == 0:
            first = dict(depth=depths.detach()[0, :, :, 0].numpy().copy(), alpha=alphas.detach()[0, :, :, 0].numpy().copy(),
                         v_depth=depths.grad[0, :, :, 0].numpy().copy(),
                         rgb_mean=renders.detach()[..., :3].mean().item())
        camera_opt.optimizer_step()
        for s in schedulers:
            s.step()
    out = {k: np.asarray(v, dtype=np.float64 if k in ("total", "Ld", "Ls", "Ln", "eT", "eR", "lr_q", "lr_t") else np.float32)
           for k, v in rows.items()}
    out["final_quat"] = camera_opt.quat_cur.detach().numpy().copy()
    out["final_t"] = camera_opt.t_cur.detach().numpy().copy()
    out.update({"first_" + k: np.asarray(v) for k, v in first.items()})
    return out


def main():
    torch.manual_seed(42)
    torch.set_num_threads(os.cpu_count() or 1)
    ref = ref_shims.install(orc.rasterization)
    fp = U.frame(U.SMALL_CAM, stride=2, seed=42)
    q0, t0 = U.offgrid_pose(fp, torch.float64)
    c2w0 = orc.pose_to_c2w(q0, t0).numpy().astype(np.float32)
    out = dict(cam=np.array([fp.W, fp.H, fp.K[0, 0], fp.K[1, 1], fp.K[0, 2], fp.K[1, 2]], dtype=np.float64),
               frame_args=np.array([2, 42]), means=fp.tar_points.astype(np.float32), scales=fp.scales.astype(np.float32),
               gt_depth=fp.src_depth.astype(np.float32), K=fp.K.astype(np.float32), c2w_init=c2w0,
               c2w_gt=fp.src_c2w.astype(np.float32), max_steps=np.array(1000))
    a = run_loop(ref, fp, c2w0, n_steps=6, max_steps=1000, depth_lambda=0.8, normal_lambda=0.0)
    out.update({"A_" + k: v for k, v in a.items()})
    b = run_loop(ref, fp, c2w0, n_steps=3, max_steps=1000, depth_lambda=0.7, normal_lambda=0.1)
    out.update({"B_" + k: v for k, v in b.items()})

    # pose module + metrics on a few seeded poses (pure reference code; kornia quaternion functions restated)
    g = torch.Generator().manual_seed(7)
    mats, quats, c2ws, eTs, eRs = [], [], [], [], []
    for i in range(6):
        axis = torch.nn.functional.normalize(torch.randn(3, generator=g), dim=0)
        ang = float(torch.rand((), generator=g)) * 3.0 + (0.1 if i else 0.0)
        Kx = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
        R = torch.eye(3) + torch.sin(torch.tensor(ang)) * Kx + (1 - torch.cos(torch.tensor(ang))) * (Kx @ Kx)
        pose = torch.eye(4); pose[:3, :3] = R; pose[:3, 3] = torch.randn(3, generator=g)
        cam = ref["model"].CameraOptModule_quat_tans(pose.clone())
        with torch.no_grad():
            cam.quat_cur.mul_(1.0 + 0.3 * i)                # un-normalised quaternion: forward() must normalise
        c2w = cam().detach()
        other = torch.eye(4); other[:3, :3] = R @ (torch.eye(3) + 0.01 * Kx @ Kx + 0.1 * Kx).qr()[0]
        other[:3, 3] = pose[:3, 3] + 0.01 * torch.randn(3, generator=g)
        mats.append(pose.numpy()); quats.append(cam.quat_cur.detach().numpy().copy()); c2ws.append(c2w.numpy())
        eTs.append(ref["eval_utils"].calculate_translation_error(c2w, other))
        eRs.append(ref["eval_utils"].calculate_rotation_error(c2w, other))
        out.setdefault("P_other", []).append(other.numpy())
    out.update(P_pose=np.asarray(mats), P_quat=np.asarray(quats), P_c2w=np.asarray(c2ws), P_eT=np.asarray(eTs),
               P_eR=np.asarray(eRs), P_other=np.asarray(out["P_other"]))
    path = os.path.join(HERE, "ref_loop_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
    print("A total:", a["total"], "\nA eT:", a["eT"], "\nB total:", b["total"], "B Ln:", b["Ln"])


if __name__ == "__main__":
    main()
