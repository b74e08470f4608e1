"""ctypes binding of libmonogs_b200.so (include/monogs_b200.h).

The library is the product; there is no CPU or eager fallback.  Importing this module on a machine where
the shared object has not been built raises, and every call maps a non-zero status to RuntimeError with
the library's own message (the reference raises RuntimeError through AT_ERROR,
submodules/diff-gaussian-rasterization/rasterize_points.cu:58-65).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmonogs_b200.so")

GRAD_MEANS3D, GRAD_COLOR, GRAD_OPACITY, GRAD_COV, GRAD_TAU, GRAD_MEANS2D = 1, 2, 4, 8, 16, 32
GRAD_ALL = 63
LOSS_TRACK_L1, LOSS_TRACK_HUBER, LOSS_MAP_L1 = 0, 1, 2


class MgsView(C.Structure):
    _fields_ = [
        ("image_height", C.c_int32), ("image_width", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("sh_degree", C.c_int32), ("sh_coeffs", C.c_int32), ("prefiltered", C.c_int32),
        ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("projmatrix_raw", C.c_void_p), ("campos", C.c_void_p),
    ]


class MgsLossArgs(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("use_depth", C.c_int32),
        ("rgb_boundary_threshold", C.c_float), ("alpha", C.c_float), ("huber_delta", C.c_float),
        ("loss_scale", C.c_float),
        ("image", C.c_void_p), ("depth", C.c_void_p), ("opacity", C.c_void_p), ("gt_image", C.c_void_p),
        ("gt_depth", C.c_void_p), ("grad_mask", C.c_void_p), ("exposure_a", C.c_void_p), ("exposure_b", C.c_void_p),
    ]


class MgsAdamGroup(C.Structure):
    _fields_ = [
        ("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
        ("numel", C.c_int64), ("lr", C.c_float), ("lr_device", C.c_void_p),
    ]


class MgsTrackArgs(C.Structure):
    _fields_ = [
        ("view", MgsView), ("P", C.c_int32),
        ("means3D", C.c_void_p), ("opacities", C.c_void_p), ("shs", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p),
        ("loss_mode", C.c_int32), ("use_depth", C.c_int32), ("fast_math", C.c_int32),
        ("rgb_boundary_threshold", C.c_float), ("alpha", C.c_float), ("huber_delta", C.c_float),
        ("gt_image", C.c_void_p), ("gt_depth", C.c_void_p), ("grad_mask", C.c_void_p),
        ("T_w2c", C.c_void_p), ("params", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
        ("step_counter", C.c_void_p), ("lr", C.c_float * 4),
        ("beta1", C.c_float), ("beta2", C.c_float), ("adam_eps", C.c_float), ("converged_threshold", C.c_float),
        ("out_color", C.c_void_p), ("out_depth", C.c_void_p), ("out_opacity", C.c_void_p),
        ("radii", C.c_void_p), ("n_touched", C.c_void_p),
        ("loss", C.c_void_p), ("tau", C.c_void_p), ("dab", C.c_void_p), ("converged", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("R_cap", C.c_int64),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("lm_state", C.c_void_p), ("status_out", C.c_void_p),
        ("tile_segments", C.c_int32), ("reserved0", C.c_int32),
    ]


class MgsTrack2Args(C.Structure):
    _fields_ = [
        ("t", MgsTrackArgs), ("n_stacks", C.c_int32), ("sketch_dim", C.c_int32),
        ("sketch_indices", C.c_void_p), ("rand_weights", C.c_void_p), ("rows", C.c_void_p), ("lm_state", C.c_void_p),
        ("min_lambda", C.c_float), ("max_lambda", C.c_float), ("increase_factor", C.c_float), ("decrease_factor", C.c_float),
        ("converged_threshold", C.c_float),
    ]


class MgsGaussians(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("sh_coeffs", C.c_int32), ("sh_degree", C.c_int32), ("isotropic", C.c_int32),
        ("xyz", C.c_void_p), ("f_dc", C.c_void_p), ("f_rest", C.c_void_p), ("opacity_raw", C.c_void_p),
        ("scaling_raw", C.c_void_p), ("rotation_raw", C.c_void_p),
    ]


class MgsMapArgs(C.Structure):
    _fields_ = [
        ("g", MgsGaussians),
        ("grad_xyz", C.c_void_p), ("grad_f_dc", C.c_void_p), ("grad_f_rest", C.c_void_p), ("grad_opacity", C.c_void_p),
        ("grad_scaling", C.c_void_p), ("grad_rotation", C.c_void_p),
        ("xyz_gradient_accum", C.c_void_p), ("denom", C.c_void_p), ("max_radii2D", C.c_void_p),
        ("loss_total", C.c_void_p), ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
    ]


class MgsMapView(C.Structure):
    _fields_ = [
        ("view", MgsView),
        ("gt_image", C.c_void_p), ("gt_depth", C.c_void_p), ("exposure_a", C.c_void_p), ("exposure_b", C.c_void_p),
        ("rgb_boundary_threshold", C.c_float), ("alpha", C.c_float), ("update_stats", C.c_int32), ("fast_math", C.c_int32), ("slot", C.c_int32), ("concurrent_backward", C.c_int32),
        ("out_color", C.c_void_p), ("out_depth", C.c_void_p), ("out_opacity", C.c_void_p),
        ("radii", C.c_void_p), ("n_touched", C.c_void_p),
        ("tau", C.c_void_p), ("dab", C.c_void_p), ("loss", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("R_cap", C.c_int64),
        ("tile_segments", C.c_int32), ("reserved0", C.c_int32),
    ]


class MgsPoseSlot(C.Structure):
    _fields_ = [
        ("T_w2c", C.c_void_p), ("params", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
        ("step", C.c_void_p), ("tau", C.c_void_p), ("dab", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("gather_out", C.c_void_p), ("lr", C.c_float * 4), ("frozen", C.c_int32),
    ]


class MgsDensifyCfg(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("n_split", C.c_int32), ("grad_threshold", C.c_float), ("min_opacity", C.c_float),
        ("extent", C.c_float), ("percent_dense", C.c_float), ("max_screen_size", C.c_float),
    ]


class MgsGaussianState(C.Structure):
    _fields_ = [
        ("param", C.c_void_p * 6), ("exp_avg", C.c_void_p * 6), ("exp_avg_sq", C.c_void_p * 6),
        ("unique_kfIDs", C.c_void_p), ("n_obs", C.c_void_p),
    ]


PEER_MAX, PEER_MAX_SEGMENTS, PEER_HANDLE_BYTES = 8, 16, 64


class MgsPeerExchange(C.Structure):
    _fields_ = [
        ("world", C.c_int32), ("rank", C.c_int32),
        ("param_arena", C.c_void_p * PEER_MAX), ("grad_arena", C.c_void_p * PEER_MAX), ("loss_out", C.c_void_p * PEER_MAX),
        ("flags", C.c_void_p * PEER_MAX),
        ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("numel", C.c_int64), ("n_segments", C.c_int32),
        ("segment_end", C.c_int64 * PEER_MAX_SEGMENTS), ("segment_lr", C.c_float * PEER_MAX_SEGMENTS),
        ("grad_multicast", C.c_void_p), ("param_multicast", C.c_void_p), ("loss_multicast", C.c_void_p),
        ("timeout_ms", C.c_uint32),
    ]


_P, _I32, _I64, _U32, _F, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_float, C.c_size_t
_VP = C.POINTER(MgsView)

# name -> (restype, argtypes); the not-gpu tests check this table against the header's declarations
SIGNATURES = {
    "mgs_last_error_string": (C.c_char_p, []),
    "mgs_version": (C.c_int, []),
    "mgs_stream_synchronize": (C.c_int, [_P]),
    "mgs_profile_enable": (C.c_int, [_I32]),
    "mgs_profile_report": (C.c_int, [C.c_char_p, _SZ]),
    "mgs_raster_workspace_bytes": (_SZ, [_I32, _I32, _I32, _I64]),
    "mgs_raster_preprocess": (C.c_int, [_VP, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P, _P]),
    "mgs_raster_render": (C.c_int, [_VP, _I32, _P, _P, _P, _SZ, _I64, _P, _P, _P, _P, _P]),
    "mgs_raster_query": (C.c_int, [_P, C.POINTER(_I32), C.POINTER(_I32), _P]),
    "mgs_raster_forward": (C.c_int, [_VP, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I64, _P, _P, _P, _P, _P, _P, _P]),
    "mgs_raster_pose_state_bytes": (_SZ, [_I32, _I32, _I32]),
    "mgs_raster_forward_pose": (C.c_int, [_VP, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I64, _P, _P, _P, _P, _P, _SZ, _P, _P, _P]),
    "mgs_raster_backward_pose_deferred": (C.c_int, [_I32, _I32, _P, _P, _P, _P, _P, _P]),
    "mgs_raster_status_async": (C.c_int, [_P, _P, _P]),
    "mgs_raster_status_reset": (C.c_int, [_P, _P]),
    "mgs_raster_backward": (C.c_int, [_VP, _I32, _U32, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _P,
                                      _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mgs_raster_backward_scratch_bytes": (_SZ, [_I32]),
    "mgs_raster_backward_pixels": (C.c_int, [_VP, _I32, _U32, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _I32, _P, _P,
                                             _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mgs_raster_backward_pixels_scratch_bytes": (_SZ, [_I32, _I32, _I32]),
    "mgs_raster_backward_pose": (C.c_int, [_VP, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _P,
                                           _P, _P, _P]),
    "mgs_raster_backward_pose_scratch_bytes": (_SZ, [_I32, _I32, _I32]),
    "mgs_sketch_reduce": (C.c_int, [_P, _I32, _I32, _I32, _I32, _P, _P, _P]),
    "mgs_mark_visible": (C.c_int, [_I32, _P, _P, _P, _P, _P]),
    "mgs_loss_fwd_bwd": (C.c_int, [C.POINTER(MgsLossArgs), _P, _P, _P, _P, _P, _SZ, _P]),
    "mgs_loss_scratch_bytes": (_SZ, [_I32, _I32]),
    "mgs_loss_fwd_bwd_static": (C.c_int, [C.POINTER(MgsLossArgs), _P, _P, _P, _P, _P]),
    "mgs_adam_step": (C.c_int, [C.POINTER(MgsAdamGroup), _I32, _I32, _F, _F, _F, _P]),
    "mgs_adam_step_device": (C.c_int, [C.POINTER(MgsAdamGroup), _I32, _P, _F, _F, _F, _P]),
    "mgs_knn_workspace_bytes": (_SZ, [_I32]),
    "mgs_knn_dist2": (C.c_int, [_I32, _P, _P, _P, _SZ, _P]),
    "mgs_track_scratch_bytes": (_SZ, [_I32, _I32, _I32]),
    "mgs_track_iteration": (C.c_int, [C.POINTER(MgsTrackArgs), _P]),
    "mgs_track_second_order": (C.c_int, [C.POINTER(MgsTrack2Args), _P]),
    "mgs_pose_update": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _F, _P]),
    "mgs_median_scratch_bytes": (_SZ, []),
    "mgs_median": (C.c_int, [_P, _I64, _P, _P, _SZ, _P]),
    "mgs_median_depth": (C.c_int, [_P, _P, _P, _I64, _P, _P, _P, _SZ, _P]),
    "mgs_frame_grad_intensity": (C.c_int, [_P, _I32, _I32, _I32, _F, _P, _P]),
    "mgs_frame_masks": (C.c_int, [_P, _I32, _I32, _F, _F, _I32, _P, _P, _P, _P, _P, _SZ, _P]),
    "mgs_visibility_pack": (C.c_int, [_I32, _P, _P, _P]),
    "mgs_covisibility": (C.c_int, [_I32, _P, _I32, _P, _P, _P]),
    "mgs_densify_scratch_bytes": (_SZ, [_I32]),
    "mgs_densify_plan": (C.c_int, [C.POINTER(MgsGaussians), _P, _P, _P, C.POINTER(MgsDensifyCfg), _P, _SZ, _P, _P]),
    "mgs_densify_apply": (C.c_int, [C.POINTER(MgsGaussians), C.POINTER(MgsGaussianState), C.POINTER(MgsGaussianState),
                                    C.POINTER(MgsDensifyCfg), _P, _P, _SZ, _P]),
    "mgs_map_scratch_bytes": (_SZ, [_I32, _I32, _I32, _I32]),
    "mgs_map_begin": (C.c_int, [C.POINTER(MgsMapArgs), _P]),
    "mgs_map_view_fwd_bwd": (C.c_int, [C.POINTER(MgsMapArgs), C.POINTER(MgsMapView), _P]),
    "mgs_map_view_forward": (C.c_int, [C.POINTER(MgsMapArgs), C.POINTER(MgsMapView), _P]),
    "mgs_map_view_backward": (C.c_int, [C.POINTER(MgsMapArgs), C.POINTER(MgsMapView), _P]),
    "mgs_map_end": (C.c_int, [C.POINTER(MgsMapArgs), _I32, _F, _P]),
    "mgs_window_pose_step": (C.c_int, [C.POINTER(MgsPoseSlot), _I32, _P, _F, _F, _F, _P]),
    "mgs_peer_alloc": (C.c_int, [_SZ, C.POINTER(C.c_void_p)]),
    "mgs_peer_free": (C.c_int, [_P]),
    "mgs_peer_export": (C.c_int, [_P, C.c_char_p]),
    "mgs_peer_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "mgs_peer_close": (C.c_int, [_P]),
    "mgs_peer_flags_bytes": (_SZ, []),
    "mgs_peer_reduce_adam": (C.c_int, [C.POINTER(MgsPeerExchange), _I32, _U32, _F, _F, _F, _P]),
    "mgs_peer_error": (C.c_int, [_P, C.POINTER(_U32), _P]),
}

_lib = None


def lib():
    """Load the shared object once; fail loudly if it is missing (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  monogs_b200 has no CPU or eager fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            if os.environ.get("MGS_DEV_PARTIAL") and not hasattr(L, name):
                continue
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().mgs_last_error_string().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (status {rc}): {msg}")


def ptr(t):
    """data pointer of a torch tensor, or None (-> NULL) for an absent / empty optional."""
    if t is None or t.numel() == 0:
        return None
    return C.c_void_p(t.data_ptr())
