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
