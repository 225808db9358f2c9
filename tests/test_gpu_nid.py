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


def test_nid_ragged_size_and_stateless_entry_points():
    """Odd image size at pyramid level 2; dms_nid_image / dms_nid_depth on caller-owned buffers."""
    import ctypes as C
    import torch
    from densemonoslam_b200 import _lib
    from oracle import nid_ref
    rng = np.random.RandomState(5)
    H, W = 61, 83
    ik = rng.randint(0, 256, (H, W)).astype(np.uint8); iko = rng.randint(0, 256, (H, W)).astype(np.uint8); ic = rng.randint(0, 256, (H, W)).astype(np.uint8)
    dk = rng.uniform(0.4, 3.4, (H, W)).astype(np.float32); dko = rng.uniform(0.4, 3.4, (H, W)).astype(np.float32); dc = rng.uniform(0.4, 3.4, (H, W)).astype(np.float32)
    for a in (dk, dko, dc):
        a[rng.rand(H, W) < 0.1] = np.nan
    L = _lib.load()
    t = lambda a: torch.from_numpy(a).cuda()
    tik, tiko, tic, tdk, tdko, tdc = t(ik), t(iko), t(ic), t(dk), t(dko), t(dc)
    for nb_i, nb_d in ((64, 500), (16, 50), (128, 1000)):
        nbm = max(nb_i, nb_d)
        hist = torch.zeros(nbm * nbm + nbm + 520, dtype=torch.int32, device="cuda")   # DMS_NID_SCRATCH_WORDS
        out = torch.zeros(4, device="cuda")
        v = C.c_float()
        rc = L.dms_nid_image(C.c_void_p(tik.data_ptr()), C.c_void_p(tiko.data_ptr()), C.c_void_p(tdk.data_ptr()), C.c_void_p(tdko.data_ptr()),
                             C.c_void_p(tic.data_ptr()), W, W * 4, W, H, W, nb_i, C.c_void_p(hist.data_ptr()), C.c_void_p(out.data_ptr()), C.byref(v), None)
        assert rc == 0
        hi = nid_ref.hist_img(ik, iko, dk, dko, ic, nb_i)
        assert np.array_equal(hist[: nb_i * nb_i].cpu().numpy().reshape(nb_i, nb_i).astype(np.int64), hi)
        ref = float(nid_ref.nid_from_hist(hi, H * W))
        assert abs(v.value - ref) <= 1e-4 * abs(ref) + 1e-6
        rc = L.dms_nid_depth(C.c_void_p(tdk.data_ptr()), C.c_void_p(tdko.data_ptr()), C.c_void_p(tdc.data_ptr()), W * 4, W * 4, H, W, nb_d, 3500.0,
                             C.c_void_p(hist.data_ptr()), C.c_void_p(out.data_ptr()), C.byref(v), None)
        assert rc == 0
        hd = nid_ref.hist_depth(dk, dko, dc, nb_d, 3500.0)
        assert np.array_equal(hist[: nb_d * nb_d].cpu().numpy().reshape(nb_d, nb_d).astype(np.int64), hd)
        ref = float(nid_ref.nid_from_hist(hd, H * W))
        assert abs(v.value - ref) <= 1e-4 * abs(ref) + 1e-6
