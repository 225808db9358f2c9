"""N4: NID keyframing score against its CPU restatement (oracle/nid_ref.py)."""
import numpy as np
import pytest

from conftest import ROOT  # noqa: F401

pytestmark = pytest.mark.gpu


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("level,nb_img,nb_depth", [(1, 64, 500), (0, 32, 100), (2, 128, 250)])
def test_nid_score_matches_restatement(level, nb_img, nb_depth):
    import torch
    from densemonoslam_b200 import synth
    from densemonoslam_b200.nid import MutualInformation
    from densemonoslam_b200.odometry import RGBDOdometry
    from oracle import nid_ref
    from oracle import oracle as orc
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(21)
    kf = synth.render(scene, cam, np.eye(4), seed=1, max_depth=50.0)                      # key frame: active model view
    old = synth.render(scene, cam, synth.pose_from((0.01, 0.0, 0.0), (0.03, 0.0, 0.0)), seed=2, max_depth=50.0)  # inactive view
    live = synth.render(scene, cam, synth.pose_from((0.02, -0.01, 0.0), (0.05, 0.02, 0.01)), seed=3, max_depth=50.0, depth_noise=0.002)
    cutoff, lam = 3.5, 0.3
    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    g.initICP(dev(torch, live["depth"]), cutoff); g.initRGB(dev(torch, live["rgba"]))
    o.initICP(live["depth"], cutoff); o.initRGB(live["rgba"])
    mi = MutualInformation(cam.width, cam.height, nb_depth, nb_img, level)
    mi.setKeyFrame(dev(torch, kf["vmap4"]), dev(torch, kf["rgba"]), dev(torch, old["vmap4"]), dev(torch, old["rgba"]), cutoff)
    ni, nd, sc = mi.score(g, cutoff, lam)
    rni, rnd, rsc, hi, hd = nid_ref.nid_score(kf["vmap4"], kf["rgba"], old["vmap4"], old["rgba"], cutoff, o.buffer("nextImage", level),
                                              o.buffer("nextDepth", level), level, nb_img, nb_depth, cutoff, lam)
    # integer work: the joint histograms are bit-exact
    assert np.array_equal(mi.histogram(0).astype(np.int64), hi)
    assert np.array_equal(mi.histogram(1).astype(np.int64), hd)
    assert hi.sum() > 0 and hd.sum() > 0
    # entropies: fp32 log2 and summation order differ (device log2f, tree sums): 1e-4 relative
    assert abs(ni - float(rni)) <= 1e-4 * abs(float(rni)) + 1e-6, (ni, rni)
    assert abs(nd - float(rnd)) <= 1e-4 * abs(float(rnd)) + 1e-6, (nd, rnd)
    assert abs(sc - float(rsc)) <= 1e-4 * abs(float(rsc)) + 1e-6
    assert 0.0 < ni <= 1.0 and 0.0 < nd <= 1.0
    # the same frame against itself is closer (smaller distance) than against the moved one
    mi.setKeyFrame(dev(torch, live["vmap4"]), dev(torch, live["rgba"]), dev(torch, live["vmap4"]), dev(torch, live["rgba"]), cutoff)
    ni_self, _, _ = mi.score(g, cutoff, lam)
    assert ni_self < ni
