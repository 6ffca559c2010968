"""Where does the rendered alpha of the 640x480 frame differ from the oracle?  (diagnostic, GPU box)"""
import sys, os, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gsb_testutil as U
from oracle import gsplat_oracle as orc
from gsplatloc_b200 import synthetic as syn, PoseTracker, TrackerConfig
fp = syn.make_frame_pair(syn.TUM, stride=2, seed=42)
q0, t0 = U.offgrid_pose(fp, torch.float64)
c2w0 = orc.pose_to_c2w(q0, t0).numpy()
means, quats, scales, opac = U.map_tensors(fp, torch.float32, "cuda")
trk = PoseTracker(fp.W, fp.H, 1, TrackerConfig(early_stop=False))
trk.set_map(means, quats, scales, opac); trk.set_observation(torch.tensor(fp.src_depth).cuda(), fp.K); trk.set_poses(c2w0, fp.src_c2w)
trk.forward()
d_gpu, a_gpu = trk.render()
a_gpu = a_gpu[0].cpu().numpy().astype(np.float64); d_gpu = d_gpu[0].cpu().numpy().astype(np.float64)
for dt in (torch.float64, torch.float32):
    m, qq, s, o = U.map_tensors(fp, dt)
    view = torch.linalg.inv(torch.tensor(c2w0, dtype=dt))[None]
    with torch.no_grad():
        rc, ra, _ = orc.rasterization(m, qq, s, o, None, view, torch.tensor(fp.K, dtype=dt)[None], fp.W, fp.H, render_mode="ED")
    a = ra[0, :, :, 0].double().numpy(); d = rc[0, :, :, 0].double().numpy()
    e = np.abs(a_gpu - a)
    print(dt, "alpha rel_err", U.rel_err(a_gpu, a), "depth rel_err", U.rel_err(d_gpu, d), "max abs", e.max(), "pixels > 1e-4:", int((e > 1e-4).sum()),
          "> 1e-5:", int((e > 1e-5).sum()), "of", e.size, "p99", np.percentile(e, 99), "median", np.median(e))
    bad = np.argwhere(e > 1e-3)[:8]
    for y, x in bad: print("   ", y, x, a_gpu[y, x], a[y, x])
