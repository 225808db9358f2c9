"""Python mirror of the result ring (include/dmsodom/ring_c.h): per-frame tracker results written device-to-device into the
memory of the consumer's GPU (peer stores over NVLink, CUDA IPC across processes)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .odometry import DmsError

HANDLE_BYTES = 64


class ResultRing:
    def __init__(self, n_sensors, slots_per_sensor, device=0, handle: bytes | None = None):
        """handle=None: consumer side, allocates the ring on `device`.  handle=bytes: producer side in another process, maps it."""
        self._L = _lib.load()
        self.n_sensors, self.slots, self.owner = int(n_sensors), int(slots_per_sensor), handle is None
        h = C.c_void_p()
        if handle is None:
            rc = self._L.dms_ring_create(C.byref(h), int(device), self.n_sensors, self.slots)
        else:
            buf = (C.c_uint8 * HANDLE_BYTES).from_buffer_copy(bytes(handle))
            rc = self._L.dms_ring_open(C.byref(h), int(device), buf, self.n_sensors, self.slots)
        if rc != 0:
            raise DmsError(f"result ring: create/open failed ({rc})")
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.dms_ring_destroy(self._h)
            self._h = None

    def export(self) -> bytes:
        buf = (C.c_uint8 * HANDLE_BYTES)()
        rc = self._L.dms_ring_export(self._h, buf)
        if rc != 0:
            raise DmsError(f"dms_ring_export failed ({rc})")
        return bytes(buf)

    def attach(self, sequence_tracker, sensor):
        rc = self._L.dms_sequence_attach_ring(sequence_tracker._h, self._h, int(sensor))
        if rc != 0:
            raise DmsError(f"dms_sequence_attach_ring failed ({rc})")
        sequence_tracker._ring = self   # keep the ring alive as long as the runner publishes into it

    def clear(self):
        rc = self._L.dms_ring_clear(self._h, None)
        if rc != 0:
            raise DmsError(f"dms_ring_clear failed ({rc})")

    def complete(self, sensor, first_frame, n_frames) -> int:
        n = C.c_int()
        rc = self._L.dms_ring_poll(self._h, int(sensor), int(first_frame), int(n_frames), C.byref(n), None)
        if rc != 0:
            raise DmsError(f"dms_ring_poll failed ({rc})")
        return n.value

    def read(self, sensor, frame):
        rec = _lib.DmsRingRecord()
        rc = self._L.dms_ring_read(self._h, int(sensor), int(frame), C.byref(rec), None)
        if rc != 0:
            raise DmsError(f"record ({sensor}, {frame}) is not complete ({rc})")
        return {"sensor": rec.sensor, "frame": rec.frame, "R": np.array(rec.R[:], np.float32).reshape(3, 3), "t": np.array(rec.t[:], np.float32),
                "stats": np.array(rec.stats[:], np.float32), "iters": list(rec.iters), "so3_iters": rec.so3_iters,
                "A": np.array(rec.A[:]).reshape(6, 6), "b": np.array(rec.b[:])}
