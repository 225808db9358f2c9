"""Python mirror of the NID keyframing entry points (include/dmsodom/nid_c.h; MutualInformation.cpp:154-212,
ElasticFusion.cpp:639-677) over libdmsodom.so.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .odometry import DmsError


class MutualInformation:
    """MutualInformation(num_bins_depth, num_bins_img, num_pyr_levels) + the KeyFrame it scores against."""

    def __init__(self, width, height, num_bins_depth=500, num_bins_img=64, pyr_level=1, device=0):
        self._L = _lib.load()
        self.nb_img, self.nb_depth, self.level = int(num_bins_img), int(num_bins_depth), int(pyr_level)
        h = C.c_void_p()
        rc = self._L.dms_nid_create(C.byref(h), int(device), int(width), int(height), self.level, self.nb_img, self.nb_depth)
        if rc != 0:
            raise DmsError(f"dms_nid_create failed ({rc})")
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.dms_nid_destroy(self._h)
            self._h = None

    def setKeyFrame(self, vertex4, rgba, old_vertex4, old_rgba, depthCutoff, stream=None):
        st = C.c_void_p(stream.cuda_stream) if stream is not None else None
        rc = self._L.dms_nid_set_keyframe(self._h, C.c_void_p(vertex4.data_ptr()), C.c_void_p(rgba.data_ptr()), C.c_void_p(old_vertex4.data_ptr()),
                                          C.c_void_p(old_rgba.data_ptr()), float(depthCutoff), st)
        if rc != 0:
            raise DmsError(f"dms_nid_set_keyframe failed ({rc})")

    def score(self, tracker, max_depth_m, depth_lambda, stream=None):
        """(nid_img, nid_depth, nid_score) against tracker.nextImg(level) / nextD(level)  (EF.cpp:651-662)."""
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        st = C.c_void_p(stream.cuda_stream) if stream is not None else None
        rc = self._L.dms_nid_score(self._h, tracker._h, float(max_depth_m), float(depth_lambda), C.byref(a), C.byref(b), C.byref(c), st)
        if rc != 0:
            raise DmsError(f"dms_nid_score failed ({rc})")
        return a.value, b.value, c.value

    def histogram(self, which):
        nb = self.nb_depth if which else self.nb_img
        out = np.zeros((nb, nb), np.uint32)
        rc = self._L.dms_nid_get_histogram(self._h, int(which), C.c_void_p(out.ctypes.data), out.size, None)
        if rc != 0:
            raise DmsError(f"dms_nid_get_histogram failed ({rc})")
        return out
