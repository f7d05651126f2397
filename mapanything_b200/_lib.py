"""ctypes binding of libmab200.so (include/mab200.h).  No CPU fallback: a missing library raises."""
import ctypes as C
import os

from . import build as _build

c_float_p = C.POINTER(C.c_float)
c_ubyte_p = C.POINTER(C.c_ubyte)


class MabConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("patch", "enc_dim", "enc_depth", "enc_heads", "enc_pos_grid",
                                       "aat_dim", "aat_depth", "aat_heads", "aat_tap0", "aat_tap1",
                                       "head_chunk", "aat_ref_view_pe", "aat_entropy_scaling", "aat_tap2", "aat_nonref_view_pe", "aat_all_global",
                                       "head_tf32", "enc_chunk", "enc_lanes")]


class MabInputs(C.Structure):
    _fields_ = [("views", C.c_int), ("H", C.c_int), ("W", C.c_int), ("images", C.c_void_p),
                ("ray_dirs", C.c_void_p), ("depth_along_ray", C.c_void_p), ("cam_quats", C.c_void_p),
                ("cam_trans", C.c_void_p), ("ray_mask", C.c_void_p), ("depth_mask", C.c_void_p),
                ("cam_mask", C.c_void_p), ("depth_metric_mask", C.c_void_p),
                ("pose_metric_mask", C.c_void_p), ("pose_trans_norm", C.c_float)]


class MabOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("pts3d", "pts3d_cam", "ray_directions", "depth_along_ray",
                                          "conf", "mask_logits", "non_ambiguous_mask", "pose",
                                          "metric_scale", "logits6")]


class MabPostOutputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("img_no_norm", "depth_z", "intrinsics", "camera_poses",
                                          "cam_trans", "cam_quats", "mask")]


class MabPostOptions(C.Structure):
    _fields_ = [("apply_mask", C.c_int), ("mask_edges", C.c_int),
                ("edge_normal_threshold_deg", C.c_float), ("edge_depth_threshold", C.c_float),
                ("apply_confidence_mask", C.c_int), ("confidence_percentile", C.c_float)]


class MabImagePlan(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("in_w", "in_h", "resize_w", "resize_h", "filter", "crop_left", "crop_top",
                                       "out_w", "out_h")]


# every symbol include/mab200.h declares: (name, restype, argtypes)
_vp, _i, _f = C.c_void_p, C.c_int, C.c_float
SYMBOLS = {
    "mab_last_error": (C.c_char_p, []),
    "mab_version": (_i, []),
    "mab_create": (_i, [C.POINTER(MabConfig), _i, C.POINTER(_vp)]),
    "mab_destroy": (_i, [_vp]),
    "mab_set_weight": (_i, [_vp, C.c_char_p, _vp, _i, _i, C.POINTER(C.c_int64)]),
    "mab_finalize_weights": (_i, [_vp, _vp]),
    "mab_set_view_pe_rows": (_i, [_vp, C.POINTER(C.c_int), _i]),
    "mab_workspace_bytes": (C.c_size_t, [_vp, _i, _i, _i, _i]),
    "mab_forward": (_i, [_vp, C.POINTER(MabInputs), C.POINTER(MabOutputs), _vp, C.c_size_t, _vp]),
    "mab_postprocess": (_i, [_vp, C.POINTER(MabInputs), C.POINTER(MabOutputs), C.POINTER(MabPostOutputs),
                             C.POINTER(MabPostOptions), _vp, C.c_size_t, _vp]),
    "mab_stage_encode": (_i, [_vp, C.POINTER(MabInputs), _i, _vp, C.c_size_t, _i, _vp]),
    "mab_stage_aat_qkv": (_i, [_vp, _i, C.POINTER(_vp), C.POINTER(_i), _vp]),
    "mab_stage_aat_attend": (_i, [_vp, _i, _vp, _i, _i, _vp]),
    "mab_stage_aat_export_kv": (_i, [_vp, _vp, _vp]),
    "mab_stage_aat_attend_kv": (_i, [_vp, _i, _vp, _i, _i, _vp]),
    "mab_stage_aat_finish": (_i, [_vp, _i, _vp]),
    "mab_stage_scale_head": (_i, [_vp, _vp, _vp]),
    "mab_stage_heads": (_i, [_vp, C.POINTER(MabInputs), C.POINTER(MabOutputs), _vp, _vp]),
    "mab_set_comm": (_i, [_vp, _vp, _i, _i, C.c_char_p]),
    "mab_kv_gather_bytes": (C.c_size_t, [_vp, _i, _i, _i]),
    "mab_forward_sharded": (_i, [_vp, C.POINTER(MabInputs), _i, C.POINTER(MabOutputs), _vp, C.c_size_t, _vp,
                                 C.c_size_t, _i, _vp]),
    "mab_rays_from_intrinsics": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "mab_normalize_rays": (_i, [_vp, C.c_longlong, _vp, _vp]),
    "mab_depth_z_to_along_ray": (_i, [_vp, _vp, C.c_longlong, _vp, _vp]),
    "mab_stage_heads_range": (_i, [_vp, C.POINTER(MabInputs), C.POINTER(MabOutputs), _i, _i, _vp]),
    "mab_set_attention_short_max": (_i, [_i]),
    "mab_launch_count": (C.c_longlong, []),
    "mab_prof_enable": (_i, [_i]),
    "mab_prof_query": (_i, [_i, C.POINTER(C.c_longlong), C.POINTER(C.c_double), C.POINTER(C.c_double),
                            C.POINTER(C.c_double)]),
    "mab_prof_records": (C.c_longlong, [C.POINTER(_i), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                        C.POINTER(C.c_double), C.c_longlong]),
    "mab_debug_buffer": (_i, [_vp, C.c_char_p, C.POINTER(_vp), C.POINTER(C.c_longlong), C.POINTER(_i)]),
    "mab_layernorm": (_i, [_vp, _i, _i, _vp, _vp, _f, _vp, _vp, _vp]),
    "mab_gemm": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "mab_attention": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp]),
    "mab_attention_carry": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _i, _i, _vp]),
    "mab_conv3x3": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "mab_pack_conv3_weight": (_i, [_vp, _i, _i, _vp, _vp]),
    "mab_gemm_tf32": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp]),
    "mab_pack_conv3_weight_f32": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "mab_conv3x3_tf32": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i, _i, _vp, _vp]),
    "mab_resize_bilinear": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "mab_posembed_resize": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "mab_rope2d": (_i, [_vp, _i, C.c_longlong, _i, _i, C.c_longlong, _vp, _f, _f, _vp]),
    "mab_image_plan_make": (_i, [_i, _i, _i, _i, C.POINTER(MabImagePlan)]),
    "mab_preprocess_workspace_bytes": (C.c_size_t, [C.POINTER(MabImagePlan)]),
    "mab_preprocess_image": (_i, [_vp, C.POINTER(MabImagePlan), c_float_p, c_float_p, _vp, _vp, _vp, C.c_size_t, _vp]),
    "mab_recover_intrinsics": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "mab_resize_nearest_crop": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "mab_export_workspace_bytes": (C.c_size_t, [_i, _i, _i]),
    "mab_export_points": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, C.c_longlong, _vp, _vp,
                               C.c_size_t, _vp]),
    "mab_export_points_static": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, C.c_longlong, _vp, _vp]),
    "mab_edge_mask": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _f, _vp, _vp, _vp]),
}

RING_FIRST, RING_LAST = 1, 2  # include/mab200.h MAB_RING_*

_lib = None


def load(allow_build=True):
    """Load libmab200.so, building it in-tree first when it is missing and nvcc is present."""
    global _lib
    if _lib is not None:
        return _lib
    # MAB_LIB: a differently-built libmab200 (kernel experiments under profiles/); the ABI is the same
    path = os.environ.get("MAB_LIB") or _build.LIB_PATH
    if not os.path.exists(path):
        if not allow_build:
            raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _build.build()
    lib = C.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        try:
            fn = getattr(lib, name)  # AttributeError if the ABI and the header ever diverge
        except AttributeError:
            if os.environ.get("MAB_LIB"):  # an experiment build of an older tree (profiles/): bind what it has
                continue
            raise
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().mab_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libmab200 {what} failed with status {status}: {msg}")
