"""CPU checks of the boundary: the library builds, loads and exports every symbol include/*.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gsplatloc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built_lib):
    handle = ctypes.CDLL(built_lib)
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/gsplatloc_b200.h but not exported"


def test_ctypes_table_matches_header(built_lib):
    from gsplatloc_b200 import _lib
    assert sorted(_lib.EXPORTS) == _declared_symbols()
    assert _lib.lib().gsb_version() >= 100


def test_struct_layouts_match_the_header(built_lib, tmp_path):
    import subprocess
    from gsplatloc_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include "gsplatloc_b200.h"\n#include <stdio.h>\nint main(void){printf("%zu %zu\\n",'
                   'sizeof(gsb_pose_state),sizeof(gsb_tracker_config));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    a, b = map(int, subprocess.check_output([str(exe)]).split())
    assert a == ctypes.sizeof(_lib.PoseState) and b == ctypes.sizeof(_lib.TrackerConfig)


def test_no_product_import_of_the_oracle():
    pkg = os.path.join(ROOT, "gsplatloc_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert "oracle" not in open(os.path.join(dp, f)).read().lower().replace("# oracle", ""), f


def test_ops_fail_loudly_without_cuda(built_lib):
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from gsplatloc_b200 import PoseTracker, ops
    with pytest.raises(RuntimeError):
        ops.fully_fused_projection(torch.zeros(4, 3), None, torch.zeros(4, 4), torch.zeros(4, 3), torch.eye(4)[None],
                                   torch.eye(3)[None], 32, 32)
    with pytest.raises(RuntimeError):
        PoseTracker(32, 32)


def test_config_defaults_equal_the_reference():
    """the numbers of CameraConfig (src/my_gsplat/model.py:18-23), OptimizationConfig
    (src/data/base.py:22-43), GsConfig (model.py:119-133) and the scheduler / early-stop constants of
    gs_trainer_total.py:65-72,160-185 are the tracker's defaults, on both sides of the C ABI."""
    import ctypes
    from gsplatloc_b200 import TrackerConfig, _lib
    c = TrackerConfig()
    assert (c.trans_lr, c.quat_lr, c.quat_opt_reg, c.trans_opt_reg) == (1e-3, 5 * 1e-4, 1e-3, 1e-3)   # model.py:19-22
    assert (c.max_steps, c.depth_lambda, c.normal_lambda) == (1000, 0.8, 0.0)                        # base.py:23-27
    assert c.early_stop is True and c.patience == 200                                                 # base.py:33-34
    assert c.warmup == 100                                                   # gs_trainer_total.py:160 "if i > 100"
    assert (c.near_plane, c.far_plane) == (1e-2, 1e10)                                                # model.py:130-133
    assert c.eps2d == 0.3 and c.radius_clip == 0.0                     # gsplat 1.3.0 rasterization() defaults (IDX:14954)
    k = _lib.TrackerConfig()
    _lib.lib().gsb_tracker_default_config(ctypes.byref(k), 1000)
    assert abs(k.lr_quat - 5e-4) < 1e-9 and abs(k.lr_trans - 1e-3) < 1e-9             # (fp32 fields)
    assert abs(k.wd_quat - 1e-3) < 1e-9 and abs(k.wd_trans - 1e-3) < 1e-9
    assert abs(k.beta1 - 0.9) < 1e-7 and abs(k.beta2 - 0.999) < 1e-7 and abs(k.adam_eps - 1e-8) < 1e-15   # torch.optim.Adam
    assert abs(k.lr_gamma - 0.2 ** (1.0 / 1000)) < 1e-7                                   # gs_trainer_total.py:66-71
    assert abs(k.depth_lambda - 0.8) < 1e-7 and k.normal_lambda == 0.0
    assert k.early_stop == 1 and k.patience == 200 and k.warmup == 100
    c2 = TrackerConfig(max_steps=2000).to_c()
    assert abs(c2.lr_gamma - 0.2 ** (1.0 / 2000)) < 1e-7


def test_align_data_mirrors_the_reference_record():
    """src/data/base.py:81-125: field names of AlignData, ``to(device)`` over every tensor attribute."""
    import torch
    from gsplatloc_b200 import AlignData, TensorWrapper
    d = AlignData(tar_points=torch.zeros(5, 3), src_points=torch.zeros(4, 3), src_depth=torch.ones(1, 6, 8, 1),
                  tar_c2w=torch.eye(4), src_c2w=torch.eye(4), tar_nums=5, pca_factor=torch.tensor(2.0))
    assert isinstance(d, TensorWrapper) and d.to("cpu") is d
    for name in ("colors", "pixels", "tar_points", "src_points", "src_depth", "tar_c2w", "src_c2w", "tar_nums", "pca_factor"):
        assert hasattr(d, name)
    fr = d.as_frame(torch.eye(3), scales=torch.ones(5, 3))
    assert fr["depth"].shape == (6, 8) and fr["means"].shape == (5, 3) and fr["c2w_init"].shape == (4, 4)
    assert set(fr) >= {"means", "quats", "scales", "opacities", "depth", "K", "c2w_init", "c2w_gt"}
