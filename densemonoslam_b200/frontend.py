"""Python mirror of the tracking front end (include/dmsodom/frontend_c.h): the tracking part of
ElasticFusion::processFrame (ElasticFusion.cpp:99-268, 397-442) over libdmsodom.so.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .odometry import DmsError

PREDICT_TRACK, PREDICT_GLOBAL, PREDICT_INACTIVE = 0, 1, 2


def _res(r):
    return {"lastICPError": r.lastICPError, "lastICPCount": r.lastICPCount, "lastRGBError": r.lastRGBError, "lastRGBCount": r.lastRGBCount,
            "lastSO3Error": r.lastSO3Error, "lastSO3Count": r.lastSO3Count, "iters": list(r.iters), "so3_iters": r.so3_iters}


class TrackingFrontEnd:
    """`predictor(pose4x4, kind, tick)` stands in for IndexMap::combinedPredict: it returns None (view not available)
    or a dict of CUDA tensors {"vertex": [H,W,4] f32, "normal": [H,W,4] f32, "image": [H,W,4] u8}."""

    def __init__(self, width, height, cx, cy, fx, fy, predictor=None, device=0, **options):
        self._L = _lib.load()
        opt = _lib.DmsFrontendOptions()
        self._L.dms_frontend_default_options(C.byref(opt))
        for k, v in options.items():
            if not hasattr(opt, k):
                raise TypeError(f"unknown front-end option {k}")
            setattr(opt, k, v)
        self.options = opt
        self.width, self.height = int(width), int(height)
        h = C.c_void_p()
        rc = self._L.dms_frontend_create(C.byref(h), int(device), self.width, self.height, float(cx), float(cy), float(fx), float(fy), C.byref(opt))
        if rc != 0:
            raise DmsError(f"dms_frontend_create failed ({rc})")
        self._h = h
        self._keep = []       # tensors returned by the predictor stay alive until the frame is done
        self._predictor = predictor
        self._cb = _lib.PREDICT_FN(self._predict)
        self._L.dms_frontend_set_predictor(self._h, self._cb, None)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.dms_frontend_destroy(self._h)
            self._h = None

    def _predict(self, user, pose, kind, tick, stream, out):
        if self._predictor is None:
            return 1
        p = np.array([pose[i] for i in range(16)], np.float32).reshape(4, 4)
        view = self._predictor(p, int(kind), int(tick))
        if view is None:
            return 1
        self._keep.append(view)
        out[0].d_vertex = view["vertex"].data_ptr(); out[0].d_normal = view["normal"].data_ptr(); out[0].d_image = view["image"].data_ptr()
        return 0

    def set_state(self, lost, lastFrameRecovery=False):
        self._L.dms_frontend_set_state(self._h, int(lost), int(lastFrameRecovery))

    def process(self, rgba, depth_raw, inPose=None, stream=None):
        """rgba: CUDA u8 [H,W,4]; depth_raw: CUDA u16/int16 [H,W] millimetres; inPose: 4x4 prior or None."""
        self._keep = []
        r = _lib.DmsFrontendResult()
        pp = None
        if inPose is not None:
            pose = np.ascontiguousarray(inPose, np.float32).reshape(16)
            pp = pose.ctypes.data_as(C.POINTER(C.c_float))
        st = C.c_void_p(stream.cuda_stream) if stream is not None else None
        rc = self._L.dms_frontend_process(self._h, C.c_void_p(rgba.data_ptr()), 0, C.c_void_p(depth_raw.data_ptr()), 0, pp, C.byref(r), st)
        if rc != 0:
            raise DmsError(f"dms_frontend_process failed ({rc})")
        return {"tick": r.tick, "tracked": bool(r.tracked), "pose": np.array(r.pose[:], np.float32).reshape(4, 4), "shouldFillIn": bool(r.shouldFillIn),
                "denseFraction": r.denseFraction, "trackingOk": bool(r.trackingOk), "lost": bool(r.lost), "trackingCount": r.trackingCount,
                "lastFrameRecovery": bool(r.lastFrameRecovery), "weighting": r.weighting, "frameToModel": _res(r.frameToModel),
                "loopChecked": bool(r.loopChecked), "loopAccepted": bool(r.loopAccepted), "loopPose": np.array(r.loopPose[:], np.float32).reshape(4, 4),
                "loopCovDiag": np.array(r.loopCovDiag[:]), "modelToModel": _res(r.modelToModel)}


def fill_in(model_vertex, model_normal, model_image, depth_filtered, live_rgba, cam, passthrough_geom=False, passthrough_img=False, stream=None):
    """FillIn::vertex / normal / image (Shaders/FillIn.cpp:65-190) on CUDA tensors; cam = (cx, cy, fx, fy)."""
    import torch
    L = _lib.load()
    H, W = depth_filtered.shape
    ov, on, oi = torch.empty_like(model_vertex), torch.empty_like(model_normal), torch.empty_like(model_image)
    st = C.c_void_p(stream.cuda_stream) if stream is not None else None
    rc = L.dms_fill_in(C.c_void_p(model_vertex.data_ptr()), C.c_void_p(model_normal.data_ptr()), C.c_void_p(model_image.data_ptr()),
                       C.c_void_p(depth_filtered.data_ptr()), C.c_void_p(live_rgba.data_ptr()), H, W, float(cam[0]), float(cam[1]), float(cam[2]),
                       float(cam[3]), int(passthrough_geom), int(passthrough_img), C.c_void_p(ov.data_ptr()), C.c_void_p(on.data_ptr()),
                       C.c_void_p(oi.data_ptr()), st)
    if rc != 0:
        raise DmsError(f"dms_fill_in failed ({rc})")
    return ov, on, oi


def dense_fraction(rgba, sample=20):
    """float(#pixels of the 1/sample image with r, g, b > 0) / #pixels (EF.cpp:84-97)."""
    import torch
    L = _lib.load()
    H, W = rgba.shape[:2]
    scratch = torch.zeros(4, dtype=torch.int32, device=rgba.device)
    f = C.c_float()
    rc = L.dms_dense_fraction(C.c_void_p(rgba.data_ptr()), H, W, int(sample), C.c_void_p(scratch.data_ptr()), C.byref(f), None)
    if rc != 0:
        raise DmsError(f"dms_dense_fraction failed ({rc})")
    return f.value


def velocity_weighting(currPose, lastPose, weightMultiplier=1.0):
    a = np.ascontiguousarray(currPose, np.float32).reshape(16); b = np.ascontiguousarray(lastPose, np.float32).reshape(16)
    fp = C.POINTER(C.c_float)
    return float(_lib.load().dms_velocity_weighting(a.ctypes.data_as(fp), b.ctypes.data_as(fp), float(weightMultiplier)))
