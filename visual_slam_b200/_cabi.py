"""ctypes binding of libvslam_b200.so (include/vslam_b200.h).

There is no CPU fallback: if the library is missing or a call returns a non-zero
status this module raises.
"""
import ctypes as C
import os

from . import build as _build

_lib = None


class VslamError(RuntimeError):
    pass


def _declare(lib):
    vp, i32, f64 = C.c_void_p, C.c_int, C.c_double
    lib.vslam_status_string.restype = C.c_char_p
    lib.vslam_status_string.argtypes = [i32]
    lib.vslam_last_cuda_error.restype = C.c_char_p
    lib.vslam_last_cuda_error.argtypes = []
    lib.vslam_abi_version.restype = i32
    lib.vslam_device_sm_count.restype = i32
    lib.vslam_hamming_knn2.argtypes = [vp, i32, vp, i32, vp, vp, vp, i32, vp]
    lib.vslam_hamming_knn2_ragged.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp, vp, i32, vp]
    lib.vslam_match_filter.argtypes = [vp, vp, vp, i32, i32, f64, i32, vp, vp, vp, i32, vp]
    lib.vslam_microbench_popc.argtypes = [i32, C.POINTER(f64), vp]
    lib.vslam_microbench_lop3.argtypes = [i32, C.POINTER(f64), vp]
    lib.vslam_hamming_set_variant.argtypes = [i32]
    pd = C.POINTER(f64)
    lib.vslam_pnp_hypotheses.argtypes = [vp, vp, vp, i32, i32, vp, i32, pd, vp, vp, vp]
    lib.vslam_pnp_score.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32, pd, f64, vp, vp]
    lib.vslam_pnp_select_refine.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, i32, pd, f64, f64, vp, i32,
                                            vp, vp, vp, vp, vp, vp]
    lib.vslam_gather_correspondences.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp, vp, vp, vp]
    lib.vslam_rodrigues.argtypes = [vp, vp, i32, vp]
    lib.vslam_rodrigues_inverse.argtypes = [vp, vp, i32, vp]
    lib.vslam_pose_camera_to_world.argtypes = [vp, vp, vp, vp, i32, vp]
    lib.vslam_pose_guess.argtypes = [vp, vp, i32, vp]
    lib.vslam_pose_guess_cam7.argtypes = [vp, vp, i32, vp]
    lib.vslam_undistort_points.argtypes = [vp, C.c_int64, pd, vp, vp]
    lib.vslam_essential_hypotheses.argtypes = [vp, vp, vp, i32, i32, vp, i32, vp, vp, vp]
    lib.vslam_essential_score.argtypes = [vp, vp, vp, i32, i32, vp, vp, i32, f64, vp, vp]
    lib.vslam_essential_select.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, i32, f64, f64, vp, vp, vp, vp,
                                           vp, vp]
    lib.vslam_essential_budget.argtypes = [vp, i32, i32, vp, vp, i32, i32, f64, vp, vp]
    lib.vslam_recover_pose.argtypes = [vp, vp, vp, vp, i32, i32, pd, f64, vp, vp, vp, vp, vp, vp]
    lib.vslam_triangulate.argtypes = [vp, vp, vp, vp, vp, i32, i32, f64, vp, vp, vp, vp, vp]
    lib.vslam_list_diff.argtypes = [vp, i32, vp, i32, i32, vp, vp, vp, vp]
    lib.vslam_features_workspace_bytes.argtypes = [i32, i32, i32, f64]
    lib.vslam_features_workspace_bytes.restype = C.c_longlong
    lib.vslam_features_set_variant.argtypes = [i32]
    lib.vslam_extract_features.argtypes = [vp, i32, i32, i32, i32, f64, f64, C.c_float, C.c_float, vp, vp,
                                           C.c_longlong, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    pp = C.POINTER(vp)
    lib.vslam_ba_create.argtypes = [pp, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.vslam_ba_destroy.argtypes = [vp]
    lib.vslam_ba_set_structure.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.vslam_ba_set_state.argtypes = [vp, vp, vp, vp, vp, C.POINTER(f64), vp]
    lib.vslam_ba_get_state.argtypes = [vp, vp, vp, vp]
    lib.vslam_ba_set_solver.argtypes = [vp, f64, i32]
    lib.vslam_ba_set_linear_solver.argtypes = [vp, i32]
    lib.vslam_ba_set_schur_variant.argtypes = [vp, i32]
    lib.vslam_ba_optimize.argtypes = [vp, i32, C.POINTER(f64), vp]
    lib.vslam_ba_optimize_batch.argtypes = [C.POINTER(vp), i32, i32, C.POINTER(f64)]
    lib.vslam_ba_linearize.argtypes = [vp, C.POINTER(f64), vp]
    lib.vslam_ba_trial.argtypes = [vp, f64, C.POINTER(f64), vp]
    lib.vslam_ba_trace_linearize.argtypes = [vp, C.POINTER(C.c_int64), C.c_int64, C.POINTER(C.c_int32), vp]
    lib.vslam_ba_time_stage.argtypes = [vp, i32, i32, vp, C.c_int64, C.POINTER(C.c_float), vp]
    lib.vslam_ba_accept.argtypes = [vp]
    lib.vslam_ba_dims.argtypes = [vp, C.POINTER(C.c_int32)]
    lib.vslam_ba_read.argtypes = [vp, i32, vp, C.c_int64, vp]
    lib.vslam_window_create.argtypes = [pp, i32, i32, i32]
    lib.vslam_window_destroy.argtypes = [vp]
    lib.vslam_window_reset.argtypes = [vp, vp, vp, i32, vp, C.POINTER(f64), vp]
    lib.vslam_window_append.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.vslam_window_motion_ba.argtypes = [vp, i32, C.POINTER(f64), vp]
    lib.vslam_window_read_cams.argtypes = [vp, vp, i32, i32, vp]
    lib.vslam_window_info.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(vp), C.POINTER(vp)]
    lib.vslam_direct_solve.argtypes = [vp, vp, vp, i32, C.POINTER(C.c_int32), vp]
    lib.vslam_pcg_solve_static.argtypes = [vp, vp, vp, i32, f64, i32, vp, vp]
    lib.vslam_pcg_solve_static_max_n.argtypes = []
    lib.vslam_pcg_solve.argtypes = [vp, vp, vp, i32, f64, i32, C.POINTER(C.c_int32), vp]


def lib():
    """Loads (building first if the sources are newer and nvcc is present) the C-ABI library."""
    global _lib
    if _lib is None:
        path = _build.LIB_PATH
        if os.path.exists(_build.NVCC):
            path = _build.build()
        if not os.path.exists(path):
            raise VslamError(
                "libvslam_b200.so is missing and nvcc is not available; "
                "run `python -m visual_slam_b200.build` (no CPU fallback exists)")
        _lib = C.CDLL(path)
        _declare(_lib)
    return _lib


def check(status):
    if status != 0:
        l = lib()
        msg = l.vslam_status_string(status).decode()
        if status == 3:
            msg += ": " + l.vslam_last_cuda_error().decode()
        raise VslamError("vslam_b200: " + msg)


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
