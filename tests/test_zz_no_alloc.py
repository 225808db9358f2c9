"""dms_icp_step_noalloc: the ICP step for callers that may not allocate device memory
(an engine with a start-up-fixed memory budget, a real-time thread that cannot take cudaMalloc's
device-wide synchronisation, a plug-in that only works on buffers it is handed).

The entry point keeps its reduction scratch in a static device array inside the library, so this
test checks the two things such a caller relies on: the result equals the handle-based
dms_odom_icp_step on the same inputs (bit for bit, the tolerance the other parity tests use for
this operator), and a second call right after the first - on a different pyramid level and pose,
so stale scratch contents would show - is just as exact.  Oversized images must be refused.

Collected last on purpose (file name): a failure here cannot hide the other tests under `-x`.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_f32_equal

pytestmark = pytest.mark.gpu


def test_noalloc_icp_step_matches_handle_entry_point_twice(synth, orc):
    import torch
    assert torch.cuda.is_available()
    from densemonoslam_b200 import _lib
    from densemonoslam_b200.odometry import RGBDOdometry, DEFAULT_ANGLE_THRESH

    L = _lib.load()
    cam = synth.CAM_VGA
    model, live, T = synth.frame_pair(cam, seed=21)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    pose = synth.pose_from((0.0, 0.01, 0.0), (0.05, 0.0, -0.02)).astype(np.float32)
    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, 0.10, DEFAULT_ANGLE_THRESH)
    g.initICPModel(dev(model["vmap4"]), dev(model["nmap4"]), 25.0, pose)
    g.initICP(dev(live["depth"]), 25.0)
    Rprev, tprev = pose[:3, :3].copy(), pose[:3, 3].copy()
    Rprev_inv = np.ascontiguousarray(np.linalg.inv(Rprev.astype(np.float64)).astype(np.float32))

    d_out = torch.zeros(C.sizeof(_lib.DmsStepSums), dtype=torch.uint8, device="cuda")  # the caller's output record
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    cases = [(0, (0.004, -0.006, 0.003), (0.008, -0.004, 0.006)), (1, (-0.003, 0.002, 0.005), (-0.006, 0.007, 0.002))]
    for level, rvec, tvec in cases:  # two calls in a row, no clean-up in between
        Tc = (pose.astype(np.float64) @ synth.pose_from(rvec, tvec)).astype(np.float32)
        Rc, tc = np.ascontiguousarray(Tc[:3, :3]), np.ascontiguousarray(Tc[:3, 3])
        ref, ref_cnt = g.icpStep(level, Rc, tc, Rprev_inv, tprev)  # existing entry point
        ptr = {}
        for name in ("vmap_curr", "nmap_curr", "vmap_g_prev", "nmap_g_prev"):
            p, pitch, rows3, cols, elem = g.buffer_ptr(name, level)
            ptr[name] = C.c_void_p(p)
        div = 1 << level
        rc = L.dms_icp_step_noalloc(ptr["vmap_curr"], ptr["nmap_curr"], ptr["vmap_g_prev"], ptr["nmap_g_prev"], pitch, rows3 // 3, cols,
                                    cam.fx / div, cam.fy / div, cam.cx / div, cam.cy / div, fp(Rc), fp(tc), fp(Rprev_inv), fp(tprev),
                                    0.10, DEFAULT_ANGLE_THRESH, C.c_void_p(d_out.data_ptr()), None)
        assert rc == 0, rc
        torch.cuda.synchronize()
        out = _lib.DmsStepSums.from_buffer_copy(d_out.cpu().numpy().tobytes())
        assert out.bad == 0
        assert out.count == ref_cnt and ref_cnt > 1000
        assert_f32_equal(np.array(out.vals[:], np.float32), ref, f"level {level}")
        # ... and the CPU oracle's icpStep on the same maps (the handle-based entry point is itself checked against it elsewhere)
        maps = {n: g.download(n, level) for n in ("vmap_curr", "nmap_curr", "vmap_g_prev", "nmap_g_prev")}
        o_vals, o_cnt = orc.icp_step(Rc, tc, maps["vmap_curr"], maps["nmap_curr"], Rprev_inv, tprev, (cam.fx / div, cam.fy / div, cam.cx / div, cam.cy / div),
                                     maps["vmap_g_prev"], maps["nmap_g_prev"], 0.10, DEFAULT_ANGLE_THRESH)
        assert out.count == o_cnt
        assert_f32_equal(np.array(out.vals[:], np.float32), o_vals, f"level {level} vs oracle")

    # beyond the stated maximum: refused before any GPU work
    rc = L.dms_icp_step_noalloc(ptr["vmap_curr"], ptr["nmap_curr"], ptr["vmap_g_prev"], ptr["nmap_g_prev"], 1 << 20, 8192, 4096,
                                1.0, 1.0, 0.0, 0.0, fp(Rc), fp(tc), fp(Rprev_inv), fp(tprev), 0.10, DEFAULT_ANGLE_THRESH,
                                C.c_void_p(d_out.data_ptr()), None)
    assert rc == -4  # DMS_ERR_TOO_LARGE
    # a frame within the pixel limit but wider than 16384 columns is refused too (the row index arithmetic is only exact up to there)
    rc = L.dms_icp_step_noalloc(ptr["vmap_curr"], ptr["nmap_curr"], ptr["vmap_g_prev"], ptr["nmap_g_prev"], 1 << 27, 2, 8000000,
                                1.0, 1.0, 0.0, 0.0, fp(Rc), fp(tc), fp(Rprev_inv), fp(tprev), 0.10, DEFAULT_ANGLE_THRESH,
                                C.c_void_p(d_out.data_ptr()), None)
    assert rc == -4
