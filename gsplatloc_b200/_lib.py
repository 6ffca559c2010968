"""ctypes binding of ``libgsplatloc_b200.so`` (the C ABI declared in ``include/gsplatloc_b200.h``).

This is the only way Python reaches the kernels: raw device pointers (``tensor.data_ptr()``), sizes and the
current CUDA stream handle.  There is no fallback — if the library is missing or a call fails, it raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgsplatloc_b200.so")

c_f = C.c_float
c_i = C.c_int
c_i64 = C.c_int64
c_p = C.c_void_p
c_sz = C.c_size_t


class TrackerConfig(C.Structure):
    """``gsb_tracker_config``."""
    _fields_ = [("lr_quat", c_f), ("lr_trans", c_f), ("wd_quat", c_f), ("wd_trans", c_f), ("beta1", c_f),
                ("beta2", c_f), ("adam_eps", c_f), ("lr_gamma", c_f), ("depth_lambda", c_f), ("normal_lambda", c_f),
                ("early_stop", c_i), ("patience", c_i), ("warmup", c_i), ("eps2d", c_f), ("near_plane", c_f),
                ("far_plane", c_f), ("radius_clip", c_f), ("normal_mode", c_i)]


class PoseState(C.Structure):
    """``gsb_pose_state``."""
    _fields_ = [("q", c_f * 4), ("t", c_f * 3), ("m_q", c_f * 4), ("v_q", c_f * 4), ("m_t", c_f * 3),
                ("v_t", c_f * 3), ("lr_q", c_f), ("lr_t", c_f), ("adam_step", c_i), ("iter", c_i), ("counter", c_i),
                ("stopped", c_i), ("best_loss", c_f), ("best_depth_loss", c_f), ("best_edge_loss", c_f),
                ("best_eT", c_f), ("best_eR", c_f), ("best_iter", c_i), ("best_q", c_f * 4), ("best_t", c_f * 3),
                ("loss", c_f), ("depth_loss", c_f), ("edge_loss", c_f), ("normal_loss", c_f), ("eT", c_f),
                ("eR", c_f), ("g_q", c_f * 4), ("g_t", c_f * 3), ("g_viewmat", c_f * 12), ("has_gt", c_i),
                ("gt_c2w", c_f * 12)]


_SIGS = {
    "gsb_last_error": (C.c_char_p, []),
    "gsb_version": (c_i, []),
    "gsb_projection_fwd": (c_i, [c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_f, c_f, c_f, c_f, c_p, c_p, c_p, c_p,
                                 c_p]),
    "gsb_projection_bwd": (c_i, [c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_f, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                 c_p, c_p]),
    "gsb_isect_count": (c_i, [c_i, c_i, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_p, C.POINTER(c_i64), c_p]),
    "gsb_isect_emit_workspace": (c_sz, [c_i, c_i, c_i64, c_i, c_i, c_i]),
    "gsb_isect_emit": (c_i, [c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_i, c_i, c_i64, c_p, c_p, c_p, c_sz,
                             c_p]),
    "gsb_isect_offset_encode": (c_i, [c_i64, c_p, c_i, c_i, c_i, c_p, c_p]),
    "gsb_rasterize_fwd": (c_i, [c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_i64, c_p, c_p, c_p,
                                c_p]),
    "gsb_rasterize_bwd": (c_i, [c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p, c_p, c_i64, c_p, c_p, c_p,
                                c_p, c_p, c_p, c_p, c_p, c_p]),
    "gsb_sh_fwd": (c_i, [c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p]),
    "gsb_sh_bwd": (c_i, [c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "gsb_depth_edge_loss": (c_i, [c_i, c_i, c_i, c_p, c_p, c_i, c_i, c_f, c_f, c_p, c_p, c_p, c_p]),
    "gsb_depth_edge_loss_static": (c_i, [c_i, c_i, c_i, c_p, c_p, c_i, c_i, c_f, c_f, c_p, c_p, c_p]),
    "gsb_loss_workspace_floats": (c_sz, [c_i, c_i, c_i]),
    "gsb_depth_edge_normal_loss": (c_i, [c_i, c_i, c_i, c_p, c_p, c_i, c_i, c_p, c_f, c_f, c_f, c_i, c_p, c_p, c_p,
                                         c_p]),
    "gsb_tracker_default_config": (None, [C.POINTER(TrackerConfig), c_i]),
    "gsb_tracker_create": (c_i, [C.POINTER(c_p), c_i, c_i, c_i, c_p]),
    "gsb_tracker_destroy": (None, [c_p]),
    "gsb_tracker_set_config": (c_i, [c_p, C.POINTER(TrackerConfig)]),
    "gsb_tracker_set_map": (c_i, [c_p, c_i, c_p, c_p, c_p, c_p, c_i]),
    "gsb_tracker_set_observation": (c_i, [c_p, c_p, c_i, c_p]),
    "gsb_tracker_set_poses": (c_i, [c_p, c_p, c_p]),
    "gsb_tracker_run": (c_i, [c_p, c_i, c_i, c_i, C.POINTER(c_i), C.POINTER(c_f)]),
    "gsb_tracker_forward": (c_i, [c_p]),
    "gsb_tracker_profile": (c_i, [c_p, c_i, c_p]),
    "gsb_tracker_get_states": (c_i, [c_p, c_p]),
    "gsb_tracker_get_render": (c_i, [c_p, c_p, c_p, c_i]),
    "gsb_tracker_get_history": (c_i, [c_p, c_i, c_p, c_i]),
    "gsb_tracker_get_counters": (c_i, [c_p, C.POINTER(c_i64), C.POINTER(c_i64), C.POINTER(c_i), C.POINTER(c_i64)]),
    "gsb_tracker_stream": (c_p, [c_p]),
    "gsb_tracker_debug_copy": (c_i, [c_p, c_p, c_p, c_p]),
    "gsb_tracker_debug_limit_bins": (c_i, [c_p, c_i]),
    "gsb_tracker_debug_loss_grads": (c_i, [c_p, c_p, c_p]),
    "gsb_tracker_set_stripe": (c_i, [c_p, c_i, c_i]),
    "gsb_tracker_stripe_segment": (c_i, [c_p, c_i]),
    "gsb_tracker_stripe_buffers": (c_i, [c_p, c_p, c_p, c_p, c_p]),
    "gsb_tracker_stripe_finish": (c_i, [c_p, c_p, c_p]),
    "gsb_tracker_stripe_exchange_buffers": (c_i, [c_p, c_p, c_p, c_p]),
    "gsb_tracker_stripe_set_peers": (c_i, [c_p, c_i, c_p, c_p, c_p]),
    "gsb_ipc_export": (c_i, [c_p, c_p]),
    "gsb_ipc_open": (c_i, [c_p, c_p]),
    "gsb_ipc_close": (c_i, [c_p]),
    "gsb_knn_sq_dists": (c_i, [c_i, c_p, c_i, c_p, C.POINTER(c_i), c_p]),
    "gsb_init_gs_scales": (c_i, [c_i, c_p, c_i, c_f, c_i, c_p, c_p]),
    "gsb_depth_to_points": (c_i, [c_i, c_i, c_p, C.POINTER(c_f), C.POINTER(c_f), c_p, c_p]),
    "gsb_transform_points": (c_i, [c_i, c_p, C.POINTER(c_f), c_p, c_p]),
    "gsb_points_median_cov": (c_i, [c_i, c_p, C.POINTER(c_f), C.POINTER(c_f), c_p]),
}

EXPORTS = tuple(_SIGS)
_lib = None


def lib():
    """Load the shared library (once).  Raises if it has not been built — there is no other code path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension is not built (run `python -m gsplatloc_b200.build` or "
                "__graft_entry__.build()).  gsplatloc_b200 has no CPU or PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().gsb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"gsplatloc_b200 {what} failed (code {rc}): {msg}")
