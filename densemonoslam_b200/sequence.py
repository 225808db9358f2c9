"""Python mirror of the sequence runner (include/dmsodom/sequence_c.h): open-loop tracking of a whole sequence with
pose priors, the trackers' host threads and streams living in the library (no interpreter lock on the path)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .odometry import DmsError


def _ptr(a):
    """device pointer of a CUDA tensor or host pointer of a (pinned) CPU tensor / numpy array"""
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return a.ctypes.data


class SequenceTracker:
    def __init__(self, width, height, cx, cy, fx, fy, device=0, **options):
        self._L = _lib.load()
        opt = _lib.DmsSeqOptions()
        self._L.dms_sequence_default_options(C.byref(opt))
        for k, v in options.items():
            if not hasattr(opt, k):
                raise TypeError(f"unknown option {k}")
            setattr(opt, k, v)
        self.options = opt
        h = C.c_void_p()
        rc = self._L.dms_sequence_create(C.byref(h), int(device), int(width), int(height), float(cx), float(cy), float(fx), float(fy), C.byref(opt))
        if rc != 0:
            raise DmsError(f"dms_sequence_create failed ({rc})")
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.dms_sequence_destroy(self._h)
            self._h = None

    @property
    def resident(self):
        return int(self._L.dms_sequence_resident(self._h))

    def host_timeline(self):
        """(first tracking kernel enqueued, last result collected, return) of the last run, microseconds since its entry"""
        us = (C.c_double * 3)()
        self._L.dms_sequence_host_timeline(self._h, us)
        return [float(v) for v in us]

    def make_frames(self, frames):
        """frames: list of dicts {"v", "n", "mi", "d", "li": tensors / arrays, "prior": 4x4}; returns the C array (keeps references)."""
        arr = (_lib.DmsSeqFrame * len(frames))()
        for a, f in zip(arr, frames):
            a.vertex4, a.normal4, a.model_rgba, a.depth, a.live_rgba = _ptr(f["v"]), _ptr(f["n"]), _ptr(f["mi"]), _ptr(f["d"]), _ptr(f["li"])
            p = np.ascontiguousarray(f["prior"], np.float32).reshape(16)
            for i in range(16):
                a.prior[i] = float(p[i])
        arr._keep = frames
        return arr

    def run(self, frames_c, first_prev_rgba, host_buffers=False):
        n = len(frames_c)
        out = (_lib.DmsSeqResult * n)()
        rc = self._L.dms_sequence_run(self._h, frames_c, n, C.c_void_p(_ptr(first_prev_rgba)), int(bool(host_buffers)), out)
        if rc != 0:
            raise DmsError(f"dms_sequence_run failed ({rc})")
        return out
