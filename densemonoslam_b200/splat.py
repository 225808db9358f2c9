"""Python mirror of the model-view synthesis entry points (include/dmsodom/splat_c.h; IndexMap::combinedPredict,
IndexMap.cpp:253-368) over libdmsodom.so.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .odometry import DmsError

ACTIVE, INACTIVE = 1, 0


def invert_pose(pose):
    a = np.ascontiguousarray(pose, np.float32).reshape(16); out = np.zeros(16, np.float32)
    fp = C.POINTER(C.c_float)
    rc = _lib.load().dms_splat_invert_pose(a.ctypes.data_as(fp), out.ctypes.data_as(fp))
    if rc != 0:
        raise DmsError(f"dms_splat_invert_pose failed ({rc})")
    return out.reshape(4, 4)


def combinedPredict(surfels, pose, cam, rows, cols, depthCutoff, confThreshold, time, timeIdx, maxTime, timeDelta, predictionType=ACTIVE, t_inv=None,
                    stream=None):
    """surfels: CUDA f32 tensor [N, 15]; cam = (cx, cy, fx, fy).  Returns (image u8 [H,W,4], vertex f32 [H,W,4], normal f32
    [H,W,4], time i32 [H,W]) CUDA tensors, zero where nothing is drawn."""
    import torch
    L = _lib.load()
    p = _lib.DmsSplatParams()
    ti = np.ascontiguousarray(invert_pose(pose) if t_inv is None else t_inv, np.float32).reshape(16)
    for i in range(16):
        p.t_inv[i] = float(ti[i])
    p.cx, p.cy, p.fx, p.fy = (float(c) for c in cam)
    p.rows, p.cols, p.maxDepth, p.confThreshold = int(rows), int(cols), float(depthCutoff), float(confThreshold)
    p.time, p.timeIdx, p.maxTime, p.timeDelta, p.active = int(time), int(timeIdx), int(maxTime), int(timeDelta), int(predictionType)
    dev = surfels.device
    zbuf = torch.empty(rows * cols, dtype=torch.int64, device=dev)
    image = torch.empty((rows, cols, 4), dtype=torch.uint8, device=dev)
    vertex = torch.empty((rows, cols, 4), dtype=torch.float32, device=dev); normal = torch.empty_like(vertex)
    tim = torch.empty((rows, cols), dtype=torch.int32, device=dev)
    st = C.c_void_p(stream.cuda_stream) if stream is not None else None
    rc = L.dms_splat_predict(C.c_void_p(surfels.data_ptr()), int(surfels.shape[0]), C.byref(p), C.c_void_p(zbuf.data_ptr()), C.c_void_p(image.data_ptr()),
                             C.c_void_p(vertex.data_ptr()), C.c_void_p(normal.data_ptr()), C.c_void_p(tim.data_ptr()), st)
    if rc != 0:
        raise DmsError(f"dms_splat_predict failed ({rc})")
    return image, vertex, normal, tim
