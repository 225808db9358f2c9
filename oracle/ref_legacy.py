"""ctypes binding of oracle/_ref/libref_legacy.so: the reference's own legacy kernels (Cuda/reduce.cu,
Cuda/cudafuncs.cu) built unmodified from /root/reference by oracle/ref_build.sh.  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libref_legacy.so")
DATATERM = np.dtype([("u0", "<i2"), ("v0", "<i2"), ("x", "<i2"), ("y", "<i2"), ("diff", "<f4"), ("valid", "u1"), ("pad", "u1", (3,))])

_lib = None


def available() -> bool:
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(SO)
        _lib.ref_device_ok.restype = C.c_int
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data)


def _f(a, n=None):
    a = np.ascontiguousarray(a, np.float32)
    return a.reshape(n) if n is not None else a


def bgr_to_intensity(rgba):
    H, W = rgba.shape[:2]; out = np.empty((H, W), np.uint8)
    lib().ref_bgr_to_intensity(_p(np.ascontiguousarray(rgba)), H, W, _p(out)); return out


def pyrdown_u8_gauss(src):
    H, W = src.shape; out = np.empty((H // 2, W // 2), np.uint8)
    lib().ref_pyrdown_u8_gauss(_p(np.ascontiguousarray(src)), H, W, _p(out)); return out


def pyrdown_f32_gauss(src):
    H, W = src.shape; out = np.empty((H // 2, W // 2), np.float32)
    lib().ref_pyrdown_f32_gauss(_p(_f(src)), H, W, _p(out)); return out


def nid_img(img_kf, img_kf_old, dmap_kf, dmap_kf_old, img_curr, nb):
    H, W = img_curr.shape
    f = lib().ref_nid_img; f.restype = C.c_float
    return float(f(_p(np.ascontiguousarray(img_kf)), _p(np.ascontiguousarray(img_kf_old)), _p(_f(dmap_kf)), _p(_f(dmap_kf_old)),
                   _p(np.ascontiguousarray(img_curr)), H, W, int(nb)))


def nid_depth(dmap_kf, dmap_kf_old, dmap_curr, nb, max_depth_mm):
    H, W = dmap_curr.shape
    f = lib().ref_nid_depth; f.restype = C.c_float
    return float(f(_p(_f(dmap_kf)), _p(_f(dmap_kf_old)), _p(_f(dmap_curr)), H, W, int(nb), C.c_float(max_depth_mm)))


def pyrdown_u16(src):
    H, W = src.shape; out = np.empty((H // 2, W // 2), np.uint16)
    lib().ref_pyrdown_u16(_p(np.ascontiguousarray(src)), H, W, _p(out)); return out


def vmap_nmap(depth, fx, fy, cx, cy, cutoff):
    H, W = depth.shape; v = np.zeros((3, H, W), np.float32); n = np.zeros((3, H, W), np.float32)
    lib().ref_vmap_nmap(_p(np.ascontiguousarray(depth)), H, W, C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), C.c_float(cutoff), _p(v), _p(n))
    return v, n


def model_maps(v4, n4, pose=None):
    H, W = v4.shape[:2]
    tot = sum(3 * (H >> i) * (W >> i) for i in range(3))
    vo = np.zeros(tot, np.float32); no = np.zeros(tot, np.float32)
    R = _f(pose[:3, :3], 9) if pose is not None else np.zeros(9, np.float32)
    t = _f(pose[:3, 3], 3) if pose is not None else np.zeros(3, np.float32)
    lib().ref_model_maps(_p(_f(v4)), _p(_f(n4)), H, W, _p(R), _p(t), int(pose is not None), _p(vo), _p(no))
    outv, outn, off = [], [], 0
    for i in range(3):
        n = 3 * (H >> i) * (W >> i)
        outv.append(vo[off:off + n].reshape(3, H >> i, W >> i)); outn.append(no[off:off + n].reshape(3, H >> i, W >> i)); off += n
    return outv, outn


def vertices_to_depth(v4, cutoff):
    H, W = v4.shape[:2]; out = np.empty((H, W), np.float32)
    lib().ref_vertices_to_depth(_p(_f(v4)), H, W, C.c_float(cutoff), _p(out)); return out


def derivatives(img):
    H, W = img.shape; dx = np.empty((H, W), np.int16); dy = np.empty((H, W), np.int16)
    lib().ref_derivatives(_p(np.ascontiguousarray(img)), H, W, _p(dx), _p(dy)); return dx, dy


def icp_step(Rcurr, tcurr, vmap_curr, nmap_curr, Rprev_inv, tprev, intr, vmap_g_prev, nmap_g_prev, distThres, angleThres, threads=128, blocks=112):
    _, H, W = vmap_curr.shape
    A = np.zeros(36, np.float32); b = np.zeros(6, np.float32); res = np.zeros(2, np.float32)
    lib().ref_icp_step(_p(_f(Rcurr, 9)), _p(_f(tcurr, 3)), _p(_f(vmap_curr)), _p(_f(nmap_curr)), _p(_f(Rprev_inv, 9)), _p(_f(tprev, 3)),
                       C.c_float(intr[0]), C.c_float(intr[1]), C.c_float(intr[2]), C.c_float(intr[3]), _p(_f(vmap_g_prev)), _p(_f(nmap_g_prev)),
                       C.c_float(distThres), C.c_float(angleThres), H, W, threads, blocks, _p(A), _p(b), _p(res))
    return A.reshape(6, 6), b, float(res[0]), int(res[1])


def rgb(minScale, dIdx, dIdy, lastDepth, nextDepth, lastImage, nextImage, maxDepthDelta, kt, krkinv, intr, sigma=None, want_step=True):
    H, W = nextImage.shape
    corr = np.zeros((H, W), DATATERM); cnt = C.c_int(); sig = C.c_int()
    A = np.zeros(36, np.float32); b = np.zeros(6, np.float32)
    lib().ref_rgb(C.c_float(minScale), _p(np.ascontiguousarray(dIdx, np.int16)), _p(np.ascontiguousarray(dIdy, np.int16)), _p(_f(lastDepth)), _p(_f(nextDepth)),
                  _p(np.ascontiguousarray(lastImage)), _p(np.ascontiguousarray(nextImage)), C.c_float(maxDepthDelta), _p(_f(kt, 3)), _p(_f(krkinv, 9)),
                  C.c_float(intr[0]), C.c_float(intr[1]), C.c_float(intr[2]), C.c_float(intr[3]),
                  C.c_float(float("nan") if sigma is None else sigma), H, W, 256, 336, 128, 112, _p(corr), C.byref(cnt), C.byref(sig),
                  _p(A) if want_step else None, _p(b) if want_step else None)
    return corr, cnt.value, sig.value, A.reshape(6, 6), b


def so3_step(lastImage, nextImage, H_, kinv, krlr, threads=160, blocks=64):
    H, W = nextImage.shape
    A = np.zeros(9, np.float32); b = np.zeros(3, np.float32); res = np.zeros(2, np.float32)
    lib().ref_so3_step(_p(np.ascontiguousarray(lastImage)), _p(np.ascontiguousarray(nextImage)), _p(_f(H_, 9)), _p(_f(kinv, 9)), _p(_f(krlr, 9)),
                       H, W, threads, blocks, _p(A), _p(b), _p(res))
    return A.reshape(3, 3), b, float(res[0]), int(res[1])


class LegacyOdometry:
    """The reference's host loop (RGBDOdometry.cpp:118-605) over the reference's own kernels; inputs are staged
    on the device in numbered slots first (the reference receives them as GL textures)."""

    def __init__(self, width, height, cx, cy, fx, fy, distThresh=0.10, angleThresh=float(np.sin(np.float32(20.0) * np.float32(3.14159254) / np.float32(180.0)))):
        L = lib()
        L.ref_odom_create.restype = C.c_void_p
        L.ref_odom_create.argtypes = [C.c_int, C.c_int] + [C.c_float] * 6
        self._h = C.c_void_p(L.ref_odom_create(width, height, cx, cy, fx, fy, distThresh, angleThresh))
        self.out = np.zeros(6, np.float32)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_odom_destroy(self._h); self._h = None

    def upload(self, slot, v4=None, n4=None, rgba=None, depth=None):
        g = lambda a, dt: None if a is None else _p(np.ascontiguousarray(a, dt))
        lib().ref_odom_upload(self._h, slot, g(v4, np.float32), g(n4, np.float32), g(rgba, np.uint8), g(depth, np.uint16))

    def initICP(self, slot, cutoff=25.0):
        lib().ref_odom_init_icp_depth(self._h, slot, C.c_float(cutoff))

    def initICPMaps(self, slot):
        lib().ref_odom_init_icp_maps(self._h, slot)

    def initICPModel(self, slot, pose):
        lib().ref_odom_init_icp_model(self._h, slot, _p(_f(pose, 16)))

    def initRGB(self, slot):
        lib().ref_odom_init_rgb(self._h, slot)

    def initRGBModel(self, slot):
        lib().ref_odom_init_rgb_model(self._h, slot)

    def initFirstRGB(self, slot):
        lib().ref_odom_init_first_rgb(self._h, slot)

    def getIncrementalTransformation(self, trans, rot, rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap=False):
        t = _f(trans, 3).copy(); R = _f(rot, 9).copy()
        lib().ref_odom_track(self._h, _p(t), _p(R), int(rgbOnly), C.c_float(icpWeight), int(pyramid), int(fastOdom), int(so3), int(interMap), _p(self.out))
        return t, R.reshape(3, 3)
