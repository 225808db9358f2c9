"""ctypes binding of the CPU oracle (oracle/cpu_odom.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_odom.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")

CORR_DTYPE = np.dtype([("u0", "<i2"), ("v0", "<i2"), ("diff", "<i2"), ("valid", "u1"), ("pad", "u1")])


class OrcResult(C.Structure):
    _fields_ = [("lastICPError", C.c_float), ("lastICPCount", C.c_float), ("lastRGBError", C.c_float),
                ("lastRGBCount", C.c_float), ("lastSO3Error", C.c_float), ("lastSO3Count", C.c_float),
                ("A", C.c_double * 36), ("b", C.c_double * 6), ("iters", C.c_int * 3), ("so3_iters", C.c_int)]


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "cpu_odom.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["bash", os.path.join(_HERE, "build.sh")])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.orc_num_threads.restype = C.c_int
        L.orc_set_num_threads.argtypes = [C.c_int]
        L.orc_odom_create.restype = C.c_void_p
        L.orc_odom_create.argtypes = [C.c_int, C.c_int] + [C.c_float] * 6
        L.orc_odom_destroy.argtypes = [C.c_void_p]
        L.orc_odom_init_icp_depth.argtypes = [C.c_void_p, _u16p, C.c_float]
        L.orc_odom_init_icp_maps.argtypes = [C.c_void_p, _f32p, _f32p]
        L.orc_odom_init_icp_model.argtypes = [C.c_void_p, _f32p, _f32p, _f32p]
        for n in ("orc_odom_init_rgb", "orc_odom_init_rgb_model", "orc_odom_init_first_rgb"):
            getattr(L, n).argtypes = [C.c_void_p, _u8p]
        L.orc_odom_track.argtypes = [C.c_void_p, _f32p, _f32p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.POINTER(OrcResult)]
        L.orc_odom_log_size.restype = C.c_int
        L.orc_odom_log_size.argtypes = [C.c_void_p]
        L.orc_odom_log_get.argtypes = [C.c_void_p, C.c_int, _i64p, _f32p, _f32p, _f32p, _f32p]
        L.orc_odom_buffer.restype = C.c_void_p
        L.orc_odom_buffer.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_covariance.argtypes = [_f64p, _f64p]
        _lib = L
    return _lib


def num_threads() -> int:
    return lib().orc_num_threads()


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(int(n))


# ----------------------------------------------------------------------------- stage functions
def bgr_to_intensity(rgba):
    H, W = rgba.shape[:2]
    out = np.empty((H, W), np.uint8)
    lib().orc_bgr_to_intensity(C.c_void_p(np.ascontiguousarray(rgba).ctypes.data), H, W, C.c_void_p(out.ctypes.data))
    return out


def pyrdown_u8_gauss(src):
    H, W = src.shape
    out = np.empty((H // 2, W // 2), np.uint8)
    lib().orc_pyrdown_u8_gauss(C.c_void_p(np.ascontiguousarray(src).ctypes.data), H, W, C.c_void_p(out.ctypes.data))
    return out


def pyrdown_f32_gauss(src):
    H, W = src.shape
    out = np.empty((H // 2, W // 2), np.float32)
    lib().orc_pyrdown_f32_gauss(C.c_void_p(np.ascontiguousarray(src, np.float32).ctypes.data), H, W, C.c_void_p(out.ctypes.data))
    return out


def pyrdown_u16(src):
    H, W = src.shape
    out = np.empty((H // 2, W // 2), np.uint16)
    lib().orc_pyrdown_u16(C.c_void_p(np.ascontiguousarray(src).ctypes.data), H, W, C.c_void_p(out.ctypes.data))
    return out


def create_vmap(depth, fx, fy, cx, cy, cutoff, prev=None):
    H, W = depth.shape
    out = np.zeros((3, H, W), np.float32) if prev is None else np.ascontiguousarray(prev, np.float32).copy()
    lib().orc_create_vmap(C.c_void_p(np.ascontiguousarray(depth).ctypes.data), H, W, C.c_float(fx), C.c_float(fy),
                          C.c_float(cx), C.c_float(cy), C.c_float(cutoff), C.c_void_p(out.ctypes.data))
    return out


def create_nmap(vmap, prev=None):
    _, H, W = vmap.shape
    out = np.zeros((3, H, W), np.float32) if prev is None else np.ascontiguousarray(prev, np.float32).copy()
    lib().orc_create_nmap(C.c_void_p(np.ascontiguousarray(vmap).ctypes.data), H, W, C.c_void_p(out.ctypes.data))
    return out


def copy_maps(v4, n4):
    H, W = v4.shape[:2]
    v = np.empty((3, H, W), np.float32)
    n = np.empty((3, H, W), np.float32)
    lib().orc_copy_maps(C.c_void_p(np.ascontiguousarray(v4).ctypes.data), C.c_void_p(np.ascontiguousarray(n4).ctypes.data), H, W,
                        C.c_void_p(v.ctypes.data), C.c_void_p(n.ctypes.data))
    return v, n


def resize_map(m, normalize, prev=None):
    _, H, W = m.shape
    out = np.zeros((3, H // 2, W // 2), np.float32) if prev is None else np.ascontiguousarray(prev, np.float32).copy()
    lib().orc_resize_map(C.c_void_p(np.ascontiguousarray(m).ctypes.data), H, W, C.c_void_p(out.ctypes.data), int(bool(normalize)))
    return out


def transform_maps(vmap, nmap, R, t):
    v = np.ascontiguousarray(vmap, np.float32).copy()
    n = np.ascontiguousarray(nmap, np.float32).copy()
    _, H, W = v.shape
    R = np.ascontiguousarray(R, np.float32).reshape(9)
    t = np.ascontiguousarray(t, np.float32).reshape(3)
    lib().orc_transform_maps(C.c_void_p(v.ctypes.data), C.c_void_p(n.ctypes.data), H, W, C.c_void_p(R.ctypes.data), C.c_void_p(t.ctypes.data))
    return v, n


def vertices_to_depth(v4, cutoff):
    H, W = v4.shape[:2]
    out = np.empty((H, W), np.float32)
    lib().orc_vertices_to_depth(C.c_void_p(np.ascontiguousarray(v4).ctypes.data), H, W, C.c_float(cutoff), C.c_void_p(out.ctypes.data))
    return out


def derivatives(img):
    H, W = img.shape
    dx = np.empty((H, W), np.int16)
    dy = np.empty((H, W), np.int16)
    lib().orc_derivatives(C.c_void_p(np.ascontiguousarray(img).ctypes.data), H, W, C.c_void_p(dx.ctypes.data), C.c_void_p(dy.ctypes.data))
    return dx, dy


def depth_bilateral(raw, maxD):
    H, W = raw.shape
    out = np.empty((H, W), np.uint16)
    lib().orc_depth_bilateral(C.c_void_p(np.ascontiguousarray(raw, np.uint16).ctypes.data), H, W, C.c_float(maxD), C.c_void_p(out.ctypes.data))
    return out


def depth_metric(raw, maxD):
    H, W = raw.shape
    out = np.empty((H, W), np.float32)
    lib().orc_depth_metric(C.c_void_p(np.ascontiguousarray(raw, np.uint16).ctypes.data), H, W, C.c_float(maxD), C.c_void_p(out.ctypes.data))
    return out


def exp_neg(t):
    L = lib(); L.orc_exp_neg.restype = C.c_float
    return L.orc_exp_neg(C.c_float(t))


def _f(a, n=None):
    a = np.ascontiguousarray(a, np.float32)
    if n is not None:
        a = a.reshape(n)
    return a


def icp_step(Rcurr, tcurr, vmap_curr, nmap_curr, Rprev_inv, tprev, intr, vmap_g_prev, nmap_g_prev, distThres, angleThres, mask=False):
    _, H, W = vmap_curr.shape
    out = np.zeros(28, np.float32)
    inl = np.zeros(1, np.int64)
    args = [C.c_void_p(_f(Rcurr, 9).ctypes.data), C.c_void_p(_f(tcurr, 3).ctypes.data), C.c_void_p(_f(vmap_curr).ctypes.data),
            C.c_void_p(_f(nmap_curr).ctypes.data), C.c_void_p(_f(Rprev_inv, 9).ctypes.data), C.c_void_p(_f(tprev, 3).ctypes.data),
            C.c_float(intr[0]), C.c_float(intr[1]), C.c_float(intr[2]), C.c_float(intr[3]),
            C.c_void_p(_f(vmap_g_prev).ctypes.data), C.c_void_p(_f(nmap_g_prev).ctypes.data), C.c_float(distThres), C.c_float(angleThres), H, W]
    lib().orc_icp_step(*args, C.c_void_p(out.ctypes.data), C.c_void_p(inl.ctypes.data))
    if mask:
        m = np.zeros((H, W), np.uint8)
        lib().orc_icp_mask(*args, C.c_void_p(m.ctypes.data))
        return out, int(inl[0]), m
    return out, int(inl[0])


def rgb_residual(minScale, dIdx, dIdy, lastDepth, nextDepth, lastImage, nextImage, maxDepthDelta, kt, krkinv):
    H, W = nextImage.shape
    corr = np.zeros((H, W), CORR_DTYPE)
    cnt = np.zeros(1, np.int64)
    sig = np.zeros(1, np.int64)
    lib().orc_rgb_residual(C.c_float(minScale), C.c_void_p(np.ascontiguousarray(dIdx, np.int16).ctypes.data),
                           C.c_void_p(np.ascontiguousarray(dIdy, np.int16).ctypes.data), C.c_void_p(_f(lastDepth).ctypes.data),
                           C.c_void_p(_f(nextDepth).ctypes.data), C.c_void_p(np.ascontiguousarray(lastImage, np.uint8).ctypes.data),
                           C.c_void_p(np.ascontiguousarray(nextImage, np.uint8).ctypes.data), C.c_float(maxDepthDelta),
                           C.c_void_p(_f(kt, 3).ctypes.data), C.c_void_p(_f(krkinv, 9).ctypes.data), H, W,
                           C.c_void_p(corr.ctypes.data), C.c_void_p(cnt.ctypes.data), C.c_void_p(sig.ctypes.data))
    return corr, int(cnt[0]), int(sig[0])


def rgb_step(corr, sigma, lastDepth, intr, dIdx, dIdy, sobelScale=0.125):
    H, W = corr.shape
    out = np.zeros(28, np.float32)
    cnt = np.zeros(1, np.int64)
    lib().orc_rgb_step(C.c_void_p(np.ascontiguousarray(corr).ctypes.data), C.c_float(sigma), C.c_void_p(_f(lastDepth).ctypes.data),
                       C.c_float(intr[0]), C.c_float(intr[1]), C.c_float(intr[2]), C.c_float(intr[3]),
                       C.c_void_p(np.ascontiguousarray(dIdx, np.int16).ctypes.data), C.c_void_p(np.ascontiguousarray(dIdy, np.int16).ctypes.data),
                       C.c_float(sobelScale), H, W, C.c_void_p(out.ctypes.data), C.c_void_p(cnt.ctypes.data))
    return out, int(cnt[0])


def so3_step(lastImage, nextImage, imageBasis, kinv, krlr):
    H, W = nextImage.shape
    out = np.zeros(10, np.float32)
    cnt = np.zeros(1, np.int64)
    lib().orc_so3_step(C.c_void_p(np.ascontiguousarray(lastImage, np.uint8).ctypes.data), C.c_void_p(np.ascontiguousarray(nextImage, np.uint8).ctypes.data),
                       C.c_void_p(_f(imageBasis, 9).ctypes.data), C.c_void_p(_f(kinv, 9).ctypes.data), C.c_void_p(_f(krlr, 9).ctypes.data),
                       H, W, C.c_void_p(out.ctypes.data), C.c_void_p(cnt.ctypes.data))
    return out, int(cnt[0])


def covariance(A):
    A = np.ascontiguousarray(A, np.float64).reshape(36)
    out = np.zeros(36, np.float64)
    lib().orc_covariance(A, out)
    return out.reshape(6, 6)


# ----------------------------------------------------------------------------- stateful tracker
BUF = {"vmap_curr": 0, "nmap_curr": 1, "vmap_g_prev": 2, "nmap_g_prev": 3, "lastDepth": 4, "nextDepth": 5, "lastImage": 6,
       "nextImage": 7, "lastNextImage": 8, "dIdx": 9, "dIdy": 10, "depth_tmp": 11, "corr": 12}
_BUF_DT = {0: (np.float32, 3), 1: (np.float32, 3), 2: (np.float32, 3), 3: (np.float32, 3), 4: (np.float32, 1), 5: (np.float32, 1),
           6: (np.uint8, 1), 7: (np.uint8, 1), 8: (np.uint8, 1), 9: (np.int16, 1), 10: (np.int16, 1), 11: (np.uint16, 1), 12: (CORR_DTYPE, 1)}


class OracleOdometry:
    """Same call surface as RGBDOdometry (RO.h:32-153) on host numpy arrays."""

    def __init__(self, width, height, cx, cy, fx, fy, distThresh=0.10, angleThresh=float(np.sin(np.float32(20.0) * np.float32(3.14159254) / np.float32(180.0)))):
        self._h = lib().orc_odom_create(width, height, cx, cy, fx, fy, distThresh, angleThresh)
        self._destroy = lib().orc_odom_destroy   # bound now: module globals may be gone at interpreter shutdown
        self.width, self.height = width, height
        self.last = None

    def __del__(self):
        if getattr(self, "_h", None):
            self._destroy(self._h)
            self._h = None

    def initICP(self, depth_u16, depthCutoff=25.0):
        lib().orc_odom_init_icp_depth(self._h, np.ascontiguousarray(depth_u16, np.uint16), depthCutoff)

    def initICPMaps(self, v4, n4):
        lib().orc_odom_init_icp_maps(self._h, _f(v4).reshape(-1), _f(n4).reshape(-1))

    def initICPModel(self, v4, n4, pose):
        lib().orc_odom_init_icp_model(self._h, _f(v4).reshape(-1), _f(n4).reshape(-1), _f(pose, 16))

    def initRGB(self, rgba):
        lib().orc_odom_init_rgb(self._h, np.ascontiguousarray(rgba, np.uint8).reshape(-1))

    def initRGBModel(self, rgba):
        lib().orc_odom_init_rgb_model(self._h, np.ascontiguousarray(rgba, np.uint8).reshape(-1))

    def initFirstRGB(self, rgba):
        lib().orc_odom_init_first_rgb(self._h, np.ascontiguousarray(rgba, np.uint8).reshape(-1))

    def getIncrementalTransformation(self, trans, rot, rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap=False, keepLog=True):
        t = _f(trans, 3).copy()
        R = _f(rot, 9).copy()
        res = OrcResult()
        lib().orc_odom_track(self._h, t, R, int(rgbOnly), C.c_float(icpWeight), int(pyramid), int(fastOdom), int(so3), int(interMap),
                             int(keepLog), C.byref(res))
        self.last = res
        return t, R.reshape(3, 3), res

    def log(self):
        n = lib().orc_odom_log_size(self._h)
        out = []
        for i in range(n):
            ints = np.zeros(5, np.int64); icp = np.zeros(28, np.float32); rgb = np.zeros(28, np.float32)
            R = np.zeros(9, np.float32); t = np.zeros(3, np.float32)
            lib().orc_odom_log_get(self._h, i, ints, icp, rgb, R, t)
            out.append({"level": int(ints[0]), "iter": int(ints[1]), "icp_count": int(ints[2]), "rgb_count": int(ints[3]),
                        "rgb_sigma": int(ints[4]), "icp": icp, "rgb": rgb, "R": R.reshape(3, 3), "t": t})
        return out

    def buffer(self, name, level=0):
        which = BUF[name]
        r = C.c_int(); c = C.c_int()
        p = lib().orc_odom_buffer(self._h, which, level, C.byref(r), C.byref(c))
        dt, planes = _BUF_DT[which]
        n = r.value * c.value * planes
        arr = np.ctypeslib.as_array((C.c_uint8 * (n * np.dtype(dt).itemsize)).from_address(p)).view(dt)
        shape = (planes, r.value, c.value) if planes == 3 else (r.value, c.value)
        return arr.reshape(shape).copy()
