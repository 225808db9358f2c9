"""Python host mirror of the reference's `RGBDOdometry` class (Utils/RGBDOdometry.h:32-153)
over the C ABI.  Same method names, argument meaning and call-order contract; GL textures
become CUDA tensors (device path) or numpy arrays (host path, copied inside the call).

PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import math
import numpy as np

from . import _lib

BUF = {"vmap_curr": 0, "nmap_curr": 1, "vmap_g_prev": 2, "nmap_g_prev": 3, "lastDepth": 4, "nextDepth": 5, "lastImage": 6,
       "nextImage": 7, "lastNextImage": 8, "dIdx": 9, "dIdy": 10, "depth_tmp": 11, "corr": 12, "nextDepthMasked": 13}
CORR_DTYPE = np.dtype([("u0", "<i2"), ("v0", "<i2"), ("diff", "<i2"), ("valid", "u1"), ("pad", "u1"), ("z", "<f4"), ("pad2", "<u4")])
_BUF_DT = {0: np.float32, 1: np.float32, 2: np.float32, 3: np.float32, 4: np.float32, 5: np.float32, 6: np.uint8, 7: np.uint8,
           8: np.uint8, 9: np.int16, 10: np.int16, 11: np.uint16, 12: CORR_DTYPE, 13: np.float32}

# default angleThresh of the reference constructor: sin(20.f * 3.14159254f / 180.f)  (RO.h:36)
DEFAULT_ANGLE_THRESH = float(np.sin(np.float32(20.0) * np.float32(3.14159254) / np.float32(180.0)))


class DmsError(RuntimeError):
    pass


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class RGBDOdometry:
    def __init__(self, width, height, cx, cy, fx, fy, distThresh=0.10, angleThresh=DEFAULT_ANGLE_THRESH, device=0):
        self._L = _lib.load()
        h = C.c_void_p()
        rc = self._L.dms_odom_create(C.byref(h), int(device), int(width), int(height), cx, cy, fx, fy, distThresh, angleThresh)
        if rc != 0:
            raise DmsError(f"dms_odom_create failed with code {rc}")
        self._h = h
        self.width, self.height, self.device = int(width), int(height), int(device)
        self.lastICPError = 0.0
        self.lastICPCount = float(width * height)
        self.lastRGBError = 0.0
        self.lastRGBCount = float(width * height)
        self.lastSO3Error = 0.0
        self.lastSO3Count = float(width * height)
        self.lastA = np.zeros((6, 6))
        self.lastb = np.zeros(6)
        self.iters = [0, 0, 0]
        self.so3_iters = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.dms_odom_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _ck(self, rc, what):
        if rc != 0:
            msg = self._L.dms_odom_last_error(self._h)
            raise DmsError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    @staticmethod
    def _stream(stream):
        if stream is None:
            return None
        return C.c_void_p(int(getattr(stream, "cuda_stream", stream)))

    def _dev(self, t, dtype_bytes_per_px):
        """(pointer, pitch) of a CUDA tensor laid out [H, W, ...] with contiguous pixels."""
        assert t.is_cuda and t.device.index == self.device
        assert t.stride(-1) == 1 or t.dim() == 2
        pitch = t.stride(0) * t.element_size()
        return C.c_void_p(t.data_ptr()), pitch

    # ------------------------------------------------------------------ RO.h surface
    def initICP(self, *args, stream=None):
        """initICP(depth, depthCutoff) or initICP(vertices, normals, depthCutoff)  (RO.h:41-45)."""
        s = self._stream(stream)
        if len(args) == 2:
            depth, cutoff = args
            if _is_torch(depth):
                p, pitch = self._dev(depth, 2)
                self._ck(self._L.dms_odom_init_icp_depth(self._h, p, pitch, cutoff, s), "initICP(depth)")
            else:
                d = np.ascontiguousarray(depth, np.uint16)
                self._ck(self._L.dms_odom_init_icp_depth_host(self._h, C.c_void_p(d.ctypes.data), 0, cutoff, s), "initICP(depth)")
        else:
            v, n, cutoff = args
            if _is_torch(v):
                pv, pitch = self._dev(v, 16)
                pn, _ = self._dev(n, 16)
                self._ck(self._L.dms_odom_init_icp_maps(self._h, pv, pn, pitch, cutoff, s), "initICP(V,N)")
            else:
                v = np.ascontiguousarray(v, np.float32); n = np.ascontiguousarray(n, np.float32)
                self._ck(self._L.dms_odom_init_icp_maps_host(self._h, C.c_void_p(v.ctypes.data), C.c_void_p(n.ctypes.data), 0, cutoff, s), "initICP(V,N)")

    def initICPModel(self, vertices, normals, depthCutoff, modelPose, stream=None):
        s = self._stream(stream)
        pose = np.ascontiguousarray(modelPose, np.float32).reshape(16)
        if _is_torch(vertices):
            pv, pitch = self._dev(vertices, 16)
            pn, _ = self._dev(normals, 16)
            self._ck(self._L.dms_odom_init_icp_model(self._h, pv, pn, pitch, depthCutoff, _fptr(pose), s), "initICPModel")
        else:
            v = np.ascontiguousarray(vertices, np.float32); n = np.ascontiguousarray(normals, np.float32)
            self._ck(self._L.dms_odom_init_icp_model_host(self._h, C.c_void_p(v.ctypes.data), C.c_void_p(n.ctypes.data), 0, depthCutoff,
                                                          _fptr(pose), s), "initICPModel")

    def _rgb(self, rgb, dev_fn, host_fn, what, stream):
        s = self._stream(stream)
        if _is_torch(rgb):
            p, pitch = self._dev(rgb, 4)
            self._ck(dev_fn(self._h, p, pitch, s), what)
        else:
            a = np.ascontiguousarray(rgb, np.uint8)
            self._ck(host_fn(self._h, C.c_void_p(a.ctypes.data), 0, s), what)

    def initRGB(self, rgb, stream=None):
        self._rgb(rgb, self._L.dms_odom_init_rgb, self._L.dms_odom_init_rgb_host, "initRGB", stream)

    def initRGBModel(self, rgb, stream=None):
        self._rgb(rgb, self._L.dms_odom_init_rgb_model, self._L.dms_odom_init_rgb_model_host, "initRGBModel", stream)

    def initFirstRGB(self, rgb, stream=None):
        self._rgb(rgb, self._L.dms_odom_init_first_rgb, self._L.dms_odom_init_first_rgb_host, "initFirstRGB", stream)

    def initFirstRGBFrom(self, src, stream=None):
        """initFirstRGB with the image of `src`'s last initRGB, copied from its pyramid on the device (dms_odom_init_first_rgb_from)."""
        self._ck(self._L.dms_odom_init_first_rgb_from(self._h, src._h, self._stream(stream)), "initFirstRGBFrom")

    def _take(self, res):
        self.lastICPError, self.lastICPCount = res.lastICPError, res.lastICPCount
        self.lastRGBError, self.lastRGBCount = res.lastRGBError, res.lastRGBCount
        self.lastSO3Error, self.lastSO3Count = res.lastSO3Error, res.lastSO3Count
        self.lastA = np.array(res.lastA[:], np.float64).reshape(6, 6)
        self.lastb = np.array(res.lastb[:], np.float64)
        self.iters = list(res.iters)
        self.so3_iters = res.so3_iters

    def getIncrementalTransformation(self, trans, rot, rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap=False, stream=None):
        """trans [3], rot [3,3] row-major: previous pose in, refined pose returned (RO.h:53-57)."""
        t = np.ascontiguousarray(trans, np.float32).reshape(3).copy()
        R = np.ascontiguousarray(rot, np.float32).reshape(9).copy()
        res = _lib.DmsResult()
        rc = self._L.dms_odom_track(self._h, _fptr(t), _fptr(R), int(bool(rgbOnly)), float(icpWeight), int(bool(pyramid)),
                                    int(bool(fastOdom)), int(bool(so3)), int(bool(interMap)), self._stream(stream), C.byref(res))
        self._ck(rc, "getIncrementalTransformation")
        self._take(res)
        return t, R.reshape(3, 3)

    def trackAsync(self, trans, rot, rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap=False, stream=None):
        t = np.ascontiguousarray(trans, np.float32).reshape(3)
        R = np.ascontiguousarray(rot, np.float32).reshape(9)
        rc = self._L.dms_odom_track_async(self._h, _fptr(t), _fptr(R), int(bool(rgbOnly)), float(icpWeight), int(bool(pyramid)),
                                          int(bool(fastOdom)), int(bool(so3)), int(bool(interMap)), self._stream(stream))
        self._ck(rc, "track_async")

    def trackCollect(self, stream=None):
        t = np.zeros(3, np.float32); R = np.zeros(9, np.float32)
        res = _lib.DmsResult()
        self._ck(self._L.dms_odom_track_collect(self._h, _fptr(t), _fptr(R), self._stream(stream), C.byref(res)), "track_collect")
        self._take(res)
        return t, R.reshape(3, 3)

    def getICPResidual(self, trans_curr, rot_curr, trans_prev, rot_prev, stream=None):
        tc = np.ascontiguousarray(trans_curr, np.float32).reshape(3); Rc = np.ascontiguousarray(rot_curr, np.float32).reshape(9)
        tp = np.ascontiguousarray(trans_prev, np.float32).reshape(3); Rp = np.ascontiguousarray(rot_prev, np.float32).reshape(9)
        err = C.c_float(); cnt = C.c_int()
        self._ck(self._L.dms_odom_icp_residual(self._h, _fptr(tc), _fptr(Rc), _fptr(tp), _fptr(Rp), C.byref(err), C.byref(cnt),
                                               self._stream(stream)), "getICPResidual")
        return err.value, cnt.value

    def getCovariance(self):
        out = (C.c_double * 36)()
        self._ck(self._L.dms_odom_get_covariance(self._h, out), "getCovariance")
        return np.array(out[:], np.float64).reshape(6, 6)

    # accessors RO.h:79-90 (downloaded copies; `buffer_ptr` gives the device view)
    def nextImg(self, level=0):
        return self.download("nextImage", level)

    def lastImg(self, level=0):
        return self.download("lastImage", level)

    def nextD(self, level=0):
        return self.download("nextDepth", level)

    def lastVmap(self, level=0):
        return self.download("vmap_g_prev", level)

    def buffer_ptr(self, name, level=0):
        p = C.c_void_p(); pitch = C.c_size_t(); rows = C.c_int(); cols = C.c_int(); elem = C.c_int()
        self._ck(self._L.dms_odom_get_buffer(self._h, BUF[name], level, C.byref(p), C.byref(pitch), C.byref(rows), C.byref(cols),
                                             C.byref(elem)), "get_buffer")
        return p.value, pitch.value, rows.value, cols.value, elem.value

    def download(self, name, level=0, stream=None):
        which = BUF[name]
        _, _, rows, cols, elem = self.buffer_ptr(name, level)
        out = np.empty(rows * cols * elem, np.uint8)
        self._ck(self._L.dms_odom_download(self._h, which, level, C.c_void_p(out.ctypes.data), out.nbytes, self._stream(stream)), "download")
        a = out.view(_BUF_DT[which])
        if which <= 3:
            return a.reshape(3, rows // 3, cols)
        return a.reshape(rows, cols)

    def set_trace(self, enabled=True):
        self._ck(self._L.dms_odom_set_trace(self._h, int(enabled)), "set_trace")

    def trace(self):
        recs = (_lib.DmsIterRecord * 160)()
        n = C.c_int()
        self._ck(self._L.dms_odom_get_trace(self._h, recs, 160, C.byref(n)), "get_trace")
        out = []
        for i in range(n.value):
            r = recs[i]
            out.append({"level": r.level, "iter": r.iter, "icp_count": r.icp_count, "rgb_count": r.rgb_count, "rgb_sigma": r.rgb_sigma,
                        "icp": np.array(r.icp[:], np.float32), "rgb": np.array(r.rgb[:], np.float32),
                        "R": np.array(r.R[:], np.float32).reshape(3, 3), "t": np.array(r.t[:], np.float32)})
        return out

    def timeline(self):
        buf = (C.c_uint64 * 2048)()
        n = C.c_int()
        self._ck(self._L.dms_odom_get_timeline(self._h, buf, 1024, C.byref(n)), "get_timeline")
        return [(int(buf[2 * i]), int(buf[2 * i + 1])) for i in range(n.value) if buf[2 * i + 1]]

    def set_grid(self, n_ctas=0):
        self._ck(self._L.dms_odom_set_grid(self._h, int(n_ctas)), "set_grid")

    def set_block(self, threads=0):
        """Threads per block of the tracking kernel (0 = 1024); call before set_grid."""
        self._ck(self._L.dms_odom_set_block(self._h, int(threads)), "set_block")

    def set_wait(self, mode=2):
        """0 spin (default), 1 blocking-sync event, 2 sleep for most of the expected wait then poll and yield."""
        self._ck(self._L.dms_odom_set_wait(self._h, int(mode)), "set_wait")

    def set_numerics(self, mode):
        """"ieee" (default, = the CPU oracle's arithmetic) or "legacy" (the reference build's approximate divisions / square roots)."""
        self._ck(self._L.dms_odom_set_numerics(self._h, {"ieee": 0, "legacy": 1}[mode] if isinstance(mode, str) else int(mode)), "set_numerics")

    def set_timing(self, mode=1):
        """0 = off, 1 = kernel events + per-phase counters, 2 = additionally the timeline of timeline()."""
        self._ck(self._L.dms_odom_set_timing(self._h, int(mode)), "set_timing")

    def timing(self):
        t = _lib.DmsTiming()
        self._ck(self._L.dms_odom_get_timing(self._h, C.byref(t)), "get_timing")
        # the reference's Stopwatch keys (RO.cpp:336,444,495,525)
        return {"track": t.track_ms, "so3Step": t.so3Step_ms, "computeRgbResidual": t.computeRgbResidual_ms,
                "icpStep": t.icpStep_ms, "rgbStep": t.rgbStep_ms, "prologue": t.prologue_ms, "serial_so3": t.serial_so3_ms,
                "serial_a": t.serial_a_ms, "serial_b": t.serial_b_ms, "barrier_wait": t.barrier_wait_ms}

    # ------------------------------------------------------------------ step operators (cudafuncs.cuh free functions)
    def icpStep(self, level, Rcurr, tcurr, Rprev_inv, tprev, want_mask=False, stream=None):
        Rc = np.ascontiguousarray(Rcurr, np.float32).reshape(9); tc = np.ascontiguousarray(tcurr, np.float32).reshape(3)
        Ri = np.ascontiguousarray(Rprev_inv, np.float32).reshape(9); tp = np.ascontiguousarray(tprev, np.float32).reshape(3)
        out = np.zeros(28, np.float32); cnt = C.c_int64()
        mask = None
        mp = None
        if want_mask:
            assert level == 0
            mask = np.zeros((self.height, self.width), np.uint8)
            mp = C.c_void_p(mask.ctypes.data)
        self._ck(self._L.dms_odom_icp_step(self._h, level, _fptr(Rc), _fptr(tc), _fptr(Ri), _fptr(tp), _fptr(out), C.byref(cnt), mp,
                                           self._stream(stream)), "icpStep")
        return (out, cnt.value, mask) if want_mask else (out, cnt.value)

    def computeDerivativeImages(self, stream=None):
        self._ck(self._L.dms_odom_compute_derivatives(self._h, self._stream(stream)), "computeDerivativeImages")

    def computeRgbResidual(self, level, kt, krkinv, stream=None):
        k = np.ascontiguousarray(kt, np.float32).reshape(3); m = np.ascontiguousarray(krkinv, np.float32).reshape(9)
        cnt = C.c_int64(); sig = C.c_int64()
        self._ck(self._L.dms_odom_rgb_residual(self._h, level, _fptr(k), _fptr(m), C.byref(cnt), C.byref(sig), self._stream(stream)),
                 "computeRgbResidual")
        return cnt.value, sig.value

    def rgbStep(self, level, sigma, stream=None):
        out = np.zeros(28, np.float32); cnt = C.c_int64()
        self._ck(self._L.dms_odom_rgb_step(self._h, level, float(sigma), _fptr(out), C.byref(cnt), self._stream(stream)), "rgbStep")
        return out, cnt.value

    def so3Step(self, level, imageBasis, kinv, krlr, stream=None):
        H = np.ascontiguousarray(imageBasis, np.float32).reshape(9); ki = np.ascontiguousarray(kinv, np.float32).reshape(9)
        kr = np.ascontiguousarray(krlr, np.float32).reshape(9)
        out = np.zeros(10, np.float32); cnt = C.c_int64()
        self._ck(self._L.dms_odom_so3_step(self._h, level, _fptr(H), _fptr(ki), _fptr(kr), _fptr(out), C.byref(cnt), self._stream(stream)),
                 "so3Step")
        return out, cnt.value


def device_count():
    return _lib.load().dms_device_count()


def version():
    return _lib.load().dms_version().decode()


def filterDepth(raw, maxDepth, stream=None):
    """ElasticFusion::filterDepth (EF.cpp:759-768): raw u16 mm CUDA tensor [H, W] -> bilateral-filtered u16 mm tensor."""
    import torch
    L = _lib.load()
    out = torch.empty_like(raw)
    rc = L.dms_depth_bilateral(C.c_void_p(raw.data_ptr()), raw.stride(0) * 2, raw.shape[0], raw.shape[1], float(maxDepth), C.c_void_p(out.data_ptr()),
                               out.stride(0) * 2, RGBDOdometry._stream(stream))
    if rc != 0:
        raise DmsError(f"dms_depth_bilateral failed ({rc})")
    return out


def metriciseDepth(raw, maxDepth, stream=None):
    """ElasticFusion::metriciseDepth (EF.cpp:748-757): u16 mm CUDA tensor -> f32 metres tensor."""
    import torch
    L = _lib.load()
    out = torch.empty(raw.shape, dtype=torch.float32, device=raw.device)
    rc = L.dms_depth_metric(C.c_void_p(raw.data_ptr()), raw.stride(0) * 2, raw.shape[0], raw.shape[1], float(maxDepth), C.c_void_p(out.data_ptr()),
                            out.stride(0) * 4, RGBDOdometry._stream(stream))
    if rc != 0:
        raise DmsError(f"dms_depth_metric failed ({rc})")
    return out
