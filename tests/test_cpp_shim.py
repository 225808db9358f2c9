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
