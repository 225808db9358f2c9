"""ctypes loader for the product library `lib/libdmsodom.so` (C ABI: include/dmsodom/odom_c.h).

There is no fallback: if the CUDA library is missing this raises, it never routes to a CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DMSODOM_LIB") or os.path.join(_HERE, "lib", "libdmsodom.so")  # DMSODOM_LIB: another build of the same library (A/B timing of kernel variants)
REPO_ROOT = os.path.dirname(_HERE)


class DmsResult(C.Structure):
    _fields_ = [("lastICPError", C.c_float), ("lastICPCount", C.c_float), ("lastRGBError", C.c_float),
                ("lastRGBCount", C.c_float), ("lastSO3Error", C.c_float), ("lastSO3Count", C.c_float),
                ("lastA", C.c_double * 36), ("lastb", C.c_double * 6), ("iters", C.c_int * 3), ("so3_iters", C.c_int)]


class DmsTiming(C.Structure):
    _fields_ = [("track_ms", C.c_float), ("so3Step_ms", C.c_float), ("computeRgbResidual_ms", C.c_float),
                ("icpStep_ms", C.c_float), ("rgbStep_ms", C.c_float), ("prologue_ms", C.c_float), ("serial_so3_ms", C.c_float),
                ("serial_a_ms", C.c_float), ("serial_b_ms", C.c_float), ("barrier_wait_ms", C.c_float)]


class DmsIterRecord(C.Structure):
    _fields_ = [("level", C.c_int32), ("iter", C.c_int32), ("icp_count", C.c_int64), ("rgb_count", C.c_int64),
                ("rgb_sigma", C.c_int64), ("icp", C.c_float * 28), ("rgb", C.c_float * 28), ("R", C.c_float * 9),
                ("t", C.c_float * 3)]


class DmsStepSums(C.Structure):
    _fields_ = [("vals", C.c_float * 28), ("count", C.c_int64), ("reserved0", C.c_int64), ("reserved1", C.c_int64),
                ("bad", C.c_int32), ("reserved2", C.c_int32)]


class DmsModelView(C.Structure):
    _fields_ = [("d_vertex", C.c_void_p), ("d_normal", C.c_void_p), ("d_image", C.c_void_p)]


class DmsFrontendOptions(C.Structure):
    _fields_ = [("depthCutoff", C.c_float), ("icpWeight", C.c_float), ("rgbOnly", C.c_int), ("pyramid", C.c_int), ("fastOdom", C.c_int),
                ("so3", C.c_int), ("frameToFrameRGB", C.c_int), ("reloc", C.c_int), ("closeLoops", C.c_int), ("hybridTracking", C.c_int),
                ("weightMultiplier", C.c_float)]


class DmsFrontendResult(C.Structure):
    _fields_ = [("tick", C.c_int), ("tracked", C.c_int), ("pose", C.c_float * 16), ("shouldFillIn", C.c_int), ("denseFraction", C.c_float),
                ("trackingOk", C.c_int), ("lost", C.c_int), ("trackingCount", C.c_int), ("lastFrameRecovery", C.c_int),
                ("weighting", C.c_float), ("frameToModel", DmsResult), ("loopChecked", C.c_int), ("loopAccepted", C.c_int),
                ("loopPose", C.c_float * 16), ("loopCovDiag", C.c_double * 6), ("modelToModel", DmsResult)]


class DmsSplatParams(C.Structure):
    _fields_ = [("t_inv", C.c_float * 16), ("cx", C.c_float), ("cy", C.c_float), ("fx", C.c_float), ("fy", C.c_float), ("rows", C.c_int),
                ("cols", C.c_int), ("maxDepth", C.c_float), ("confThreshold", C.c_float), ("time", C.c_int), ("timeIdx", C.c_int),
                ("maxTime", C.c_int), ("timeDelta", C.c_int), ("active", C.c_int)]


class DmsSeqFrame(C.Structure):
    _fields_ = [("vertex4", C.c_void_p), ("normal4", C.c_void_p), ("model_rgba", C.c_void_p), ("depth", C.c_void_p), ("live_rgba", C.c_void_p),
                ("prior", C.c_float * 16)]


class DmsSeqResult(C.Structure):
    _fields_ = [("pose", C.c_float * 16), ("r", DmsResult), ("track_ms", C.c_float), ("host_ms", C.c_float)]


class DmsRingRecord(C.Structure):
    _fields_ = [("seq", C.c_uint32), ("sensor", C.c_uint32), ("frame", C.c_uint32), ("reserved", C.c_uint32), ("R", C.c_float * 9), ("t", C.c_float * 3),
                ("stats", C.c_float * 6), ("iters", C.c_int32 * 3), ("so3_iters", C.c_int32), ("A", C.c_double * 36), ("b", C.c_double * 6),
                ("pad", C.c_uint8 * 72)]


class DmsSeqOptions(C.Structure):
    _fields_ = [("n_trackers", C.c_int), ("block_threads", C.c_int), ("grid", C.c_int), ("wait_mode", C.c_int), ("rgbOnly", C.c_int),
                ("icpWeight", C.c_float), ("pyramid", C.c_int), ("fastOdom", C.c_int), ("so3", C.c_int), ("depthCutoff", C.c_float),
                ("timing", C.c_int)]


# int (*dms_predict_fn)(void* user, const float pose[16], int kind, int tick, void* stream, dms_model_view* out)
PREDICT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_float), C.c_int, C.c_int, C.c_void_p, C.POINTER(DmsModelView))

# every symbol declared in include/dmsodom/*.h (tests check the headers against this list)
_P = C.c_void_p
_F3 = C.POINTER(C.c_float)
SIGNATURES = {
    "dms_odom_create": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int] + [C.c_float] * 6),
    "dms_odom_destroy": (C.c_int, [_P]),
    "dms_odom_last_error": (C.c_char_p, [_P]),
    "dms_odom_set_grid": (C.c_int, [_P, C.c_int]),
    "dms_odom_set_block": (C.c_int, [_P, C.c_int]),
    "dms_odom_set_wait": (C.c_int, [_P, C.c_int]),
    "dms_odom_set_numerics": (C.c_int, [_P, C.c_int]),
    "dms_odom_set_timing": (C.c_int, [_P, C.c_int]),
    "dms_odom_get_timing": (C.c_int, [_P, C.POINTER(DmsTiming)]),
    "dms_odom_init_icp_depth": (C.c_int, [_P, _P, C.c_size_t, C.c_float, _P]),
    "dms_odom_init_icp_maps": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_float, _P]),
    "dms_odom_init_icp_model": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_float, _F3, _P]),
    "dms_odom_init_rgb": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_init_rgb_model": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_init_first_rgb": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_init_first_rgb_from": (C.c_int, [_P, _P, _P]),
    "dms_odom_init_icp_depth_host": (C.c_int, [_P, _P, C.c_size_t, C.c_float, _P]),
    "dms_odom_init_icp_maps_host": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_float, _P]),
    "dms_odom_init_icp_model_host": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_float, _F3, _P]),
    "dms_odom_init_rgb_host": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_init_rgb_model_host": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_init_first_rgb_host": (C.c_int, [_P, _P, C.c_size_t, _P]),
    "dms_odom_track": (C.c_int, [_P, _F3, _F3, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.POINTER(DmsResult)]),
    "dms_odom_track_async": (C.c_int, [_P, _F3, _F3, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "dms_odom_track_collect": (C.c_int, [_P, _F3, _F3, _P, C.POINTER(DmsResult)]),
    "dms_odom_icp_residual": (C.c_int, [_P, _F3, _F3, _F3, _F3, C.POINTER(C.c_float), C.POINTER(C.c_int), _P]),
    "dms_odom_get_covariance": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "dms_odom_get_buffer": (C.c_int, [_P, C.c_int, C.c_int, C.POINTER(_P), C.POINTER(C.c_size_t), C.POINTER(C.c_int),
                                      C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "dms_odom_download": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_size_t, _P]),
    "dms_odom_set_trace": (C.c_int, [_P, C.c_int]),
    "dms_odom_get_trace": (C.c_int, [_P, C.POINTER(DmsIterRecord), C.c_int, C.POINTER(C.c_int)]),
    "dms_odom_get_timeline": (C.c_int, [_P, C.POINTER(C.c_uint64), C.c_int, C.POINTER(C.c_int)]),
    "dms_odom_icp_step": (C.c_int, [_P, C.c_int, _F3, _F3, _F3, _F3, _F3, C.POINTER(C.c_int64), _P, _P]),
    "dms_odom_rgb_residual": (C.c_int, [_P, C.c_int, _F3, _F3, C.POINTER(C.c_int64), C.POINTER(C.c_int64), _P]),
    "dms_odom_rgb_step": (C.c_int, [_P, C.c_int, C.c_float, _F3, C.POINTER(C.c_int64), _P]),
    "dms_odom_so3_step": (C.c_int, [_P, C.c_int, _F3, _F3, _F3, _F3, C.POINTER(C.c_int64), _P]),
    "dms_odom_compute_derivatives": (C.c_int, [_P, _P]),
    "dms_depth_bilateral": (C.c_int, [_P, C.c_size_t, C.c_int, C.c_int, C.c_float, _P, C.c_size_t, _P]),
    "dms_depth_metric": (C.c_int, [_P, C.c_size_t, C.c_int, C.c_int, C.c_float, _P, C.c_size_t, _P]),
    "dms_icp_step_noalloc": (C.c_int, [_P, _P, _P, _P, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                       _F3, _F3, _F3, _F3, C.c_float, C.c_float, _P, _P]),
    "dms_frontend_default_options": (None, [C.POINTER(DmsFrontendOptions)]),
    "dms_frontend_create": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int] + [C.c_float] * 4 + [C.POINTER(DmsFrontendOptions)]),
    "dms_frontend_destroy": (C.c_int, [_P]),
    "dms_frontend_set_predictor": (C.c_int, [_P, PREDICT_FN, _P]),
    "dms_frontend_process": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_size_t, _F3, C.POINTER(DmsFrontendResult), _P]),
    "dms_frontend_get_tracker": (C.c_int, [_P, C.c_int, C.POINTER(_P)]),
    "dms_frontend_get_filtered_depth": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "dms_frontend_get_fill_in": (C.c_int, [_P, C.POINTER(DmsModelView)]),
    "dms_frontend_set_state": (C.c_int, [_P, C.c_int, C.c_int]),
    "dms_fill_in": (C.c_int, [_P, _P, _P, _P, _P, C.c_int, C.c_int] + [C.c_float] * 4 + [C.c_int, C.c_int, _P, _P, _P, _P]),
    "dms_dense_fraction": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, C.POINTER(C.c_float), _P]),
    "dms_velocity_weighting": (C.c_float, [_F3, _F3, C.c_float]),
    "dms_nid_create": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "dms_nid_destroy": (C.c_int, [_P]),
    "dms_nid_set_keyframe": (C.c_int, [_P, _P, _P, _P, _P, C.c_float, _P]),
    "dms_nid_score": (C.c_int, [_P, _P, C.c_float, C.c_float, _F3, _F3, _F3, _P]),
    "dms_nid_get_histogram": (C.c_int, [_P, C.c_int, _P, C.c_size_t, _P]),
    "dms_nid_image": (C.c_int, [_P, _P, _P, _P, _P, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, _P, _P, _F3, _P]),
    "dms_nid_depth": (C.c_int, [_P, _P, _P, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_float, _P, _P, _F3, _P]),
    "dms_splat_predict": (C.c_int, [_P, C.c_int, C.POINTER(DmsSplatParams), _P, _P, _P, _P, _P, _P]),
    "dms_splat_invert_pose": (C.c_int, [_F3, _F3]),
    "dms_sequence_default_options": (None, [C.POINTER(DmsSeqOptions)]),
    "dms_sequence_create": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int] + [C.c_float] * 4 + [C.POINTER(DmsSeqOptions)]),
    "dms_sequence_destroy": (C.c_int, [_P]),
    "dms_sequence_run": (C.c_int, [_P, C.POINTER(DmsSeqFrame), C.c_int, _P, C.c_int, C.POINTER(DmsSeqResult)]),
    "dms_sequence_resident": (C.c_int, [_P]),
    "dms_sequence_host_timeline": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "dms_sequence_attach_ring": (C.c_int, [_P, _P, C.c_int]),
    "dms_ring_create": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int]),
    "dms_ring_export": (C.c_int, [_P, C.POINTER(C.c_uint8)]),
    "dms_ring_open": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(C.c_uint8), C.c_int, C.c_int]),
    "dms_ring_destroy": (C.c_int, [_P]),
    "dms_ring_clear": (C.c_int, [_P, _P]),
    "dms_ring_publish": (C.c_int, [_P, _P, C.c_int, C.c_int, _P]),
    "dms_ring_poll": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), _P]),
    "dms_ring_read": (C.c_int, [_P, C.c_int, C.c_int, C.POINTER(DmsRingRecord), _P]),
    "dms_version": (C.c_char_p, []),
    "dms_device_count": (C.c_int, []),
    "dms_track_d2h_bytes": (C.c_size_t, []),
    "dms_threshold_equiv": (None, [C.c_float, C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
}

_lib = None


def load():
    """Load libdmsodom.so and bind every entry point.  Raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with densemonoslam_b200/csrc/build.sh "
                "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
