"""The C++ drop-in surface (include/dmsodom/RGBDOdometry.h, same class surface as the reference's
Utils/RGBDOdometry.h:32-153) driven the way ElasticFusion.cpp:173-201 / GPUTest.cpp:260-283 drive it, against the
Python mirror on the same inputs: results must be identical (both sit on the same C ABI)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "tests", "cpp", "_build", "shim_test")


def test_cpp_shim_matches_python_mirror(tmp_path, synth):
    import torch
    assert torch.cuda.is_available()
    if not os.path.exists(BIN):
        subprocess.check_call(["bash", os.path.join(ROOT, "tests", "cpp", "build.sh")])
    from densemonoslam_b200.odometry import RGBDOdometry
    cam = synth.CAM_VGA
    model, live, T = synth.frame_pair(cam, seed=1234)
    d = str(tmp_path)
    model["vmap4"].tofile(os.path.join(d, "vmap4.bin")); model["nmap4"].tofile(os.path.join(d, "nmap4.bin"))
    model["rgba"].tofile(os.path.join(d, "model_rgba.bin")); live["rgba"].tofile(os.path.join(d, "live_rgba.bin")); live["depth"].tofile(os.path.join(d, "depth.bin"))
    open(os.path.join(d, "meta.txt"), "w").write(f"{cam.width} {cam.height} {cam.cx} {cam.cy} {cam.fx} {cam.fy}\n")
    subprocess.check_call([BIN, d], timeout=120)
    out = [list(map(float, l.split())) for l in open(os.path.join(d, "out.txt"))]

    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    pose = np.eye(4, dtype=np.float32)
    g.initFirstRGB(dev(model["rgba"])); g.initICPModel(dev(model["vmap4"]), dev(model["nmap4"]), 25.0, pose); g.initRGBModel(dev(model["rgba"]))
    g.initICP(dev(live["depth"]), 25.0); g.initRGB(dev(live["rgba"]))
    t, R = g.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True)
    assert np.array_equal(np.float32(out[0]), t)
    assert np.array_equal(np.float32(out[1:4]), R)
    assert np.allclose(out[4], [g.lastICPError, g.lastICPCount, g.lastRGBError, g.lastRGBCount, g.lastSO3Error, g.lastSO3Count], rtol=1e-6)
    assert out[5][0] == g.lastA[0, 0] and abs(out[5][1] - g.getCovariance()[0, 0]) <= 1e-12 * abs(out[5][1])
    assert out[5][2] == float(g.nextImg(0).astype(np.uint64).sum())   # the accessor sees the swapped pyramid (RO.cpp:595-601)
    assert out[6] == [1.0, 1.0, 1.0, 1.0]  # the four Stopwatch keys GPUTest.cpp:280-283 reads exist
    assert np.max(np.abs(t - T[:3, 3])) < 5e-3


def test_cpp_front_end_mirrors_match_python(tmp_path, synth):
    """include/dmsodom/ElasticFusionFrontEnd.h (IndexMap::combinedPredict, the processFrame tracking block, MutualInformation)
    driven from C++ against the Python mirrors on the same inputs: identical results (same C ABI underneath)."""
    import torch
    from densemonoslam_b200 import splat
    from densemonoslam_b200.frontend import TrackingFrontEnd
    from densemonoslam_b200.nid import MutualInformation
    from densemonoslam_b200.odometry import RGBDOdometry
    from oracle import splat_ref as sr
    subprocess.check_call(["bash", os.path.join(ROOT, "tests", "cpp", "build.sh")])
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(8)
    c = (cam.cx, cam.cy, cam.fx, cam.fy)
    v0 = synth.render(scene, cam, np.eye(4), seed=1, max_depth=50.0, hole_frac=0.0)
    S = sr.surfels_from_view(v0["vmap4"], v0["nmap4"], v0["rgba"], np.eye(4), c, time=1)
    T1 = synth.pose_from((0.003, -0.004, 0.001), (0.008, -0.003, 0.004))
    live0 = synth.render(scene, cam, np.eye(4), seed=50, max_depth=50.0, depth_noise=0.001)
    live1 = synth.render(scene, cam, T1, seed=51, max_depth=50.0, depth_noise=0.001)
    prior1 = (T1 @ synth.pose_from((0.004, 0.003, -0.002), (0.006, -0.005, 0.004))).astype(np.float32)
    d = str(tmp_path)
    S.tofile(os.path.join(d, "surfels.bin")); prior1.tofile(os.path.join(d, "prior1.bin"))
    for i, l in enumerate((live0, live1)):
        l["rgba"].tofile(os.path.join(d, f"live{i}_rgba.bin")); l["depth"].tofile(os.path.join(d, f"live{i}_depth.bin"))
    open(os.path.join(d, "meta.txt"), "w").write(f"{cam.width} {cam.height} {cam.cx} {cam.cy} {cam.fx} {cam.fy}\n")
    subprocess.check_call([os.path.join(ROOT, "tests", "cpp", "_build", "frontend_test"), d], timeout=120)
    out = [list(map(float, l.split())) for l in open(os.path.join(d, "out_fe.txt"))]

    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dS = dev(S)

    def predictor(pose, kind, tick):
        if kind == 2:
            return None
        img, ver, nor, _ = splat.combinedPredict(dS, pose, c, cam.height, cam.width, 3.5, 5.0, tick, 0, tick, 200, splat.ACTIVE)
        return {"vertex": ver, "normal": nor, "image": img}

    fe = TrackingFrontEnd(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor, depthCutoff=3.5, closeLoops=0)
    r0 = fe.process(dev(live0["rgba"]), dev(live0["depth"]), np.eye(4, dtype=np.float32))
    r1 = fe.process(dev(live1["rgba"]), dev(live1["depth"]), prior1)
    assert out[0] == [r0["tick"], float(r0["tracked"]), r1["tick"], float(r1["tracked"])] == [1.0, 0.0, 2.0, 1.0]
    assert np.array_equal(np.float32(out[1:5]), r1["pose"])
    assert out[5][0] == r1["frameToModel"]["lastICPCount"] and out[5][1] == r1["frameToModel"]["lastRGBCount"]
    assert abs(out[5][2] - r1["weighting"]) < 1e-6 and out[5][3] == float(r1["shouldFillIn"]) and out[5][4] == float(r1["trackingOk"])
    # NID score of the tracked frame against the prediction at the tracked pose
    img, ver, nor, _ = splat.combinedPredict(dS, r1["pose"], c, cam.height, cam.width, 3.5, 5.0, 2, 0, 2, 200, splat.ACTIVE)
    mi = MutualInformation(cam.width, cam.height, 500, 64, 1)
    mi.setKeyFrame(ver, img, ver, img, 3.5)
    import ctypes as C
    from densemonoslam_b200 import _lib
    h = C.c_void_p()
    assert _lib.load().dms_frontend_get_tracker(fe._h, 0, C.byref(h)) == 0

    class _T:   # the front end's frameToModel tracker as the nid wrapper expects it
        _h = h
    ni, nd, sc = mi.score(_T, 3.5, 0.5)
    assert np.allclose(out[6], [sc, ni, nd], rtol=1e-6)
    assert 0.0 < sc < 1.0
