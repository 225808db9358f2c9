"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.

Bar (BASELINE.json north_star): integer outputs (intensity images, derivative images, inlier /
correspondence masks and counts) bit-exact; the reduced JtJ/Jtr within 1e-4 relative (they are
in fact bit-identical here because oracle and kernels share the canonical summation tree); final
pose within 1e-5 m / 1e-5 rad.
"""
import numpy as np
import pytest

from conftest import assert_f32_equal, assert_maps_equal

pytestmark = pytest.mark.gpu

RTOL_SYSTEM = 1e-4   # relative tolerance on the reduced system (north_star)
POSE_TOL = 1e-5      # metres / radians


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch


def make_pair(synth, cam, seed=1234, **kw):
    return synth.frame_pair(cam, seed=seed, **kw)


def both(orc, synth, cam, distThresh=0.10):
    from densemonoslam_b200.odometry import RGBDOdometry, DEFAULT_ANGLE_THRESH
    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, distThresh, DEFAULT_ANGLE_THRESH)
    g.set_trace(True)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, distThresh, DEFAULT_ANGLE_THRESH)
    return g, o


def to_dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def load_frame_to_model(torch, g, o, model, live, pose, device_inputs=True):
    """EF.cpp:173-185 order: initICPModel, initRGBModel, initICP(depth), initRGB."""
    cv = (lambda a: to_dev(torch, a)) if device_inputs else (lambda a: a)
    g.initICPModel(cv(model["vmap4"]), cv(model["nmap4"]), 25.0, pose)
    g.initRGBModel(cv(model["rgba"]))
    g.initICP(cv(live["depth"]), 25.0)
    g.initRGB(cv(live["rgba"]))
    o.initICPModel(model["vmap4"], model["nmap4"], pose)
    o.initRGBModel(model["rgba"])
    o.initICP(live["depth"], 25.0)
    o.initRGB(live["rgba"])


def check_pyramids(g, o, names_maps=("vmap_curr", "nmap_curr", "vmap_g_prev", "nmap_g_prev")):
    for l in range(3):
        for n in names_maps:
            assert_maps_equal(g.download(n, l), o.buffer(n, l), f"{n}[{l}]")
        for n in ("lastDepth", "nextDepth"):
            assert_f32_equal(g.download(n, l), o.buffer(n, l), f"{n}[{l}]")
        for n in ("lastImage", "nextImage", "depth_tmp"):
            assert np.array_equal(g.download(n, l), o.buffer(n, l)), f"{n}[{l}]"


def compare_track(g, o, trans, rot, flags, check_trace=True):
    tg, Rg = g.getIncrementalTransformation(trans, rot, *flags)
    to_, Ro, res = o.getIncrementalTransformation(trans, rot, *flags)
    assert g.iters == list(res.iters), (g.iters, list(res.iters))
    assert g.so3_iters == res.so3_iters
    if check_trace:
        tr_g, tr_o = g.trace(), o.log()
        assert len(tr_g) == len(tr_o)
        for a, b in zip(tr_g, tr_o):
            assert (a["level"], a["iter"]) == (b["level"], b["iter"])
            # integer work: bit-exact at every iteration
            assert a["icp_count"] == b["icp_count"], (a["level"], a["iter"], a["icp_count"], b["icp_count"])
            assert a["rgb_count"] == b["rgb_count"], (a["level"], a["iter"])
            assert a["rgb_sigma"] == b["rgb_sigma"], (a["level"], a["iter"])
            for k in ("icp", "rgb"):
                scale = max(1e-30, float(np.max(np.abs(b[k]))))
                assert np.max(np.abs(a[k] - b[k])) <= RTOL_SYSTEM * scale, (k, a["level"], a["iter"])
            assert np.max(np.abs(a["t"] - b["t"])) <= POSE_TOL
            assert np.max(np.abs(a["R"] - b["R"])) <= POSE_TOL
    # public fields
    assert g.lastICPCount == res.lastICPCount
    assert g.lastRGBCount == res.lastRGBCount
    assert g.lastSO3Count == res.lastSO3Count
    for a, b in ((g.lastICPError, res.lastICPError), (g.lastRGBError, res.lastRGBError), (g.lastSO3Error, res.lastSO3Error)):
        assert (np.isnan(a) and np.isnan(b)) or abs(a - b) <= 1e-4 * max(abs(b), 1e-12), (a, b)
    A_o = np.array(res.A[:]).reshape(6, 6); b_o = np.array(res.b[:])
    assert np.max(np.abs(g.lastA - A_o)) <= RTOL_SYSTEM * max(np.max(np.abs(A_o)), 1e-30)
    assert np.max(np.abs(g.lastb - b_o)) <= RTOL_SYSTEM * max(np.max(np.abs(b_o)), 1e-30)
    # final pose
    assert np.max(np.abs(tg - to_)) <= POSE_TOL, (tg, to_)
    assert np.max(np.abs(Rg - Ro)) <= POSE_TOL
    return tg, Rg


# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("camname", ["vga", "kitti", "ferns80x60"])
def test_pyramid_construction_bit_exact(torch_cuda, orc, synth, camname):
    cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "ferns80x60": synth.scaled_camera(synth.CAM_VGA, 8)}[camname]
    scene = synth.street_scene(3) if camname == "kitti" else None
    model, live, T = make_pair(synth, cam, seed=11, scene=scene, max_depth=60.0)
    g, o = both(orc, synth, cam)
    pose = synth.pose_from((0.02, -0.01, 0.03), (0.1, -0.2, 0.05)).astype(np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
    check_pyramids(g, o)
    for l in range(3):
        assert np.array_equal(g.download("lastNextImage", l), o.buffer("lastNextImage", l))
    # model-to-model order: the second vertex map feeds nextDepth (RO.cpp:213,246; EF.cpp:410-418)
    g.initICP(to_dev(torch_cuda, live["vmap4"]), to_dev(torch_cuda, live["nmap4"]), 25.0); o.initICPMaps(live["vmap4"], live["nmap4"])
    g.initRGB(to_dev(torch_cuda, live["rgba"])); o.initRGB(live["rgba"])
    check_pyramids(g, o)


def test_host_buffer_entry_points_match_device(torch_cuda, orc, synth, pair_small):
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)
    load_frame_to_model(torch_cuda, g, o, model, live, np.eye(4, dtype=np.float32), device_inputs=False)
    check_pyramids(g, o)


def test_derivative_images_bit_exact(torch_cuda, orc, synth, pair_vga):
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    load_frame_to_model(torch_cuda, g, o, model, live, np.eye(4, dtype=np.float32))
    g.computeDerivativeImages()
    for l in range(3):
        dx, dy = orc.derivatives(o.buffer("nextImage", l))
        assert np.array_equal(g.download("dIdx", l), dx)
        assert np.array_equal(g.download("dIdy", l), dy)


@pytest.mark.parametrize("level", [0, 1, 2])
def test_icp_step_parity(torch_cuda, orc, synth, pair_vga, level):
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    pose = synth.pose_from((0.0, 0.01, 0.0), (0.05, 0.0, -0.02)).astype(np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    Rprev, tprev = pose[:3, :3].copy(), pose[:3, 3].copy()
    Rprev_inv = np.linalg.inv(Rprev.astype(np.float64)).astype(np.float32)
    Tc = (pose.astype(np.float64) @ synth.pose_from((0.004, -0.006, 0.003), (0.008, -0.004, 0.006))).astype(np.float32)
    Rc, tc = Tc[:3, :3].copy(), Tc[:3, 3].copy()
    div = 1 << level
    intr = (np.float32(cam.fx) / div, np.float32(cam.fy) / div, np.float32(cam.cx) / div, np.float32(cam.cy) / div)
    from densemonoslam_b200.odometry import DEFAULT_ANGLE_THRESH
    args = (Rc, tc, o.buffer("vmap_curr", level), o.buffer("nmap_curr", level), Rprev_inv, tprev, intr,
            o.buffer("vmap_g_prev", level), o.buffer("nmap_g_prev", level), 0.10, DEFAULT_ANGLE_THRESH)
    if level == 0:
        ref, inl, mask = orc.icp_step(*args, mask=True)
        out, cnt, gmask = g.icpStep(level, Rc, tc, Rprev_inv, tprev, want_mask=True)
        assert np.array_equal(gmask, mask), f"{np.count_nonzero(gmask != mask)} inlier-mask pixels differ"
    else:
        ref, inl = orc.icp_step(*args)
        out, cnt = g.icpStep(level, Rc, tc, Rprev_inv, tprev)
    assert cnt == inl and inl > 1000
    assert_f32_equal(out, ref, "icp 27 products + residual")


@pytest.mark.parametrize("level", [0, 1, 2])
def test_rgb_residual_and_step_parity(torch_cuda, orc, synth, pair_vga, level):
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    load_frame_to_model(torch_cuda, g, o, model, live, np.eye(4, dtype=np.float32))
    g.computeDerivativeImages()
    div = 1 << level
    K = np.array([[cam.fx / div, 0, cam.cx / div], [0, cam.fy / div, cam.cy / div], [0, 0, 1]], np.float64)
    Rt = synth.pose_from((0.003, -0.004, 0.002), (0.006, -0.003, 0.004))
    krkinv = (K @ Rt[:3, :3] @ np.linalg.inv(K)).astype(np.float32)
    kt = (K @ Rt[:3, 3]).astype(np.float32)
    dx, dy = orc.derivatives(o.buffer("nextImage", level))
    minScale = {0: 1600.0, 1: 576.0, 2: 64.0}[level]
    corr, cnt, sig = orc.rgb_residual(minScale, dx, dy, o.buffer("lastDepth", level), o.buffer("nextDepth", level),
                                      o.buffer("lastImage", level), o.buffer("nextImage", level), 0.07, kt, krkinv)
    gcnt, gsig = g.computeRgbResidual(level, kt, krkinv)
    assert (gcnt, gsig) == (cnt, sig) and cnt > 100
    gcorr = g.download("corr", level)
    assert np.array_equal(gcorr["valid"], corr["valid"])
    for f in ("u0", "v0", "diff"):
        assert np.array_equal(gcorr[f], corr[f]), f
    intr = (np.float32(cam.fx) / div, np.float32(cam.fy) / div, np.float32(cam.cx) / div, np.float32(cam.cy) / div)
    for sigma in (float(np.float32(np.sqrt(float(cnt)))), -1.0):
        ref, c2 = orc.rgb_step(corr, sigma, o.buffer("lastDepth", level), intr, dx, dy)
        out, c3 = g.rgbStep(level, sigma)
        assert c2 == c3 == cnt
        assert_f32_equal(out, ref, f"rgb products sigma={sigma}")


def test_so3_step_parity(torch_cuda, orc, synth, pair_vga):
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    load_frame_to_model(torch_cuda, g, o, model, live, np.eye(4, dtype=np.float32))
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
    level = 2
    K = np.array([[cam.fx / 4, 0, cam.cx / 4], [0, cam.fy / 4, cam.cy / 4], [0, 0, 1]], np.float64)
    R = synth.rodrigues((0.01, -0.02, 0.005))
    H = (K @ R @ np.linalg.inv(K)).astype(np.float32); kinv = np.linalg.inv(K).astype(np.float32); krlr = (K @ R).astype(np.float32)
    ref, cnt = orc.so3_step(o.buffer("lastNextImage", level), o.buffer("nextImage", level), H, kinv, krlr)
    out, gcnt = g.so3Step(level, H, kinv, krlr)
    assert gcnt == cnt and cnt > 1000
    assert_f32_equal(out, ref, "so3 products")


# usage matrix of SURVEY.md 8(b): (rgbOnly, icpWeight, pyramid, fastOdom, so3, interMap)
FLAG_ROWS = {
    "frameToModel": (False, 10.0, True, False, True, False),
    "frameToModel_fastOdom_nopyr": (False, 10.0, False, True, True, False),
    "modelToModel": (False, 10.0, True, False, False, False),
    "rgbOnly": (True, 10.0, True, False, True, False),
    "icpOnly_w100": (False, 100.0, True, False, False, False),
    "gputest": (False, 10.0, False, False, True, False),
    "referenceFrame": (False, 10.0, True, False, True, True),   # ReferenceFrame.h:82-90: ICP + RGB, so3, interMap => 50/50/50
}


@pytest.mark.parametrize("row", sorted(FLAG_ROWS))
def test_full_track_parity_vga(torch_cuda, orc, synth, pair_vga, row):
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
    t, R = compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS[row])
    if row == "frameToModel":  # the tracker recovers the rendered motion
        assert np.max(np.abs(t - T[:3, 3])) < 5e-3
        assert np.max(np.abs(R - T[:3, :3])) < 2e-3
    # second frame on the same handles: so3 image swap (RO.cpp:595-601) and buffer reuse
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS[row])
    for l in range(3):
        assert np.array_equal(g.download("nextImage", l), o.buffer("nextImage", l))
        assert np.array_equal(g.download("lastNextImage", l), o.buffer("lastNextImage", l))


def test_reference_frame_call_shape(torch_cuda, orc, synth, pair_vga):
    """ReferenceFrame::m_rgbd (ReferenceFrame.h:82-90; usage matrix row 4 of SURVEY 8b): full resolution, initICPModel ->
    initICP(V, N) -> initRGBModel -> initRGB, so BOTH depth pyramids come from the second vertex map (RO.cpp:213,240,246),
    then ICP + RGB with SO(3) and interMap = 150 Gauss-Newton iterations, every one compared."""
    cam, model, live, T = pair_vga
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    cv = lambda a: to_dev(torch_cuda, a)
    g.initFirstRGB(cv(model["rgba"])); o.initFirstRGB(model["rgba"])
    g.initICPModel(cv(model["vmap4"]), cv(model["nmap4"]), 25.0, pose); o.initICPModel(model["vmap4"], model["nmap4"], pose)
    g.initICP(cv(live["vmap4"]), cv(live["nmap4"]), 25.0); o.initICPMaps(live["vmap4"], live["nmap4"])
    g.initRGBModel(cv(model["rgba"])); o.initRGBModel(model["rgba"])
    g.initRGB(cv(live["rgba"])); o.initRGB(live["rgba"])
    check_pyramids(g, o)
    compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["referenceFrame"])
    assert g.iters == [50, 50, 50] and len(g.trace()) == 150


def mat3_inv_f32(m):
    """Eigen's fixed-size 3x3 .inverse() by cofactors in fp32, in the operation order the library and the oracle use."""
    m = [np.float32(x) for x in np.asarray(m, np.float32).reshape(9)]
    c00 = m[4] * m[8] - m[5] * m[7]; c01 = m[5] * m[6] - m[3] * m[8]; c02 = m[3] * m[7] - m[4] * m[6]
    det = m[0] * c00 + m[1] * c01 + m[2] * c02
    i = np.float32(1) / det
    return np.array([c00 * i, (m[2] * m[7] - m[1] * m[8]) * i, (m[1] * m[5] - m[2] * m[4]) * i,
                     c01 * i, (m[0] * m[8] - m[2] * m[6]) * i, (m[2] * m[3] - m[0] * m[5]) * i,
                     c02 * i, (m[1] * m[6] - m[0] * m[7]) * i, (m[0] * m[4] - m[1] * m[3]) * i], np.float32).reshape(3, 3)


def test_icp_residual_against_oracle(torch_cuda, orc, synth, pair_small):
    """getICPResidual (RO.cpp:612-643): level-0 residual-only ICP with the reference's odd argument shape -- rot_curr.inverse()
    as Rprev_inv and trans_prev unchanged (RO.cpp:623-625) -- against the oracle's icpStep on the same arguments."""
    from densemonoslam_b200.odometry import DEFAULT_ANGLE_THRESH
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)
    pose = synth.pose_from((0.01, -0.02, 0.015), (0.03, 0.01, -0.02)).astype(np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    Tc = (pose.astype(np.float64) @ synth.pose_from((0.004, -0.006, 0.003), (0.008, -0.004, 0.006))).astype(np.float32)
    Rc, tc = np.ascontiguousarray(Tc[:3, :3]), np.ascontiguousarray(Tc[:3, 3])
    err, cnt = g.getICPResidual(tc, Rc, pose[:3, 3], pose[:3, :3])
    p, inl = orc.icp_step(Rc, tc, o.buffer("vmap_curr", 0), o.buffer("nmap_curr", 0), mat3_inv_f32(Rc), pose[:3, 3], (cam.fx, cam.fy, cam.cx, cam.cy),
                          o.buffer("vmap_g_prev", 0), o.buffer("nmap_g_prev", 0), 0.10, DEFAULT_ANGLE_THRESH)
    assert cnt == inl and inl > 1000
    assert np.float32(err) == np.sqrt(np.float32(p[27])) / np.float32(inl)


def test_full_track_parity_kitti_shape(torch_cuda, orc, synth):
    cam = synth.CAM_KITTI
    frames = synth.kitti_sequence(cam, 2, seed=2021)
    g, o = both(orc, synth, cam)
    g.initFirstRGB(to_dev(torch_cuda, frames[0]["live"]["rgba"])); o.initFirstRGB(frames[0]["live"]["rgba"])
    for f in frames:
        prior = f["prior"]
        load_frame_to_model(torch_cuda, g, o, f["model"], f["live"], prior)
        compare_track(g, o, prior[:3, 3].copy(), prior[:3, :3].copy(), FLAG_ROWS["frameToModel"])


def test_ferns_shape_intermap(torch_cuda, orc, synth):
    """Ferns::rgbd / ReferenceFrame usage: W/8 x H/8, ICP only (w=100) and interMap 50/50/50 (Ferns.cpp:363-383)."""
    cam = synth.scaled_camera(synth.CAM_VGA, 8)
    model, live, T = synth.frame_pair(cam, seed=5, rvec=(0.004, -0.006, 0.003), tvec=(0.008, -0.004, 0.006))
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    cv = lambda a: to_dev(torch_cuda, a)
    g.initICPModel(cv(model["vmap4"]), cv(model["nmap4"]), 25.0, pose); o.initICPModel(model["vmap4"], model["nmap4"], pose)
    g.initICP(cv(live["vmap4"]), cv(live["nmap4"]), 25.0); o.initICPMaps(live["vmap4"], live["nmap4"])
    compare_track(g, o, pose[:3, 3], pose[:3, :3], (False, 100.0, False, False, False, False))
    compare_track(g, o, pose[:3, 3], pose[:3, :3], (False, 100.0, True, False, False, True))


def test_identity_motion_and_covariance(torch_cuda, orc, synth, pair_small):
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    m2 = dict(model); m2["depth"] = model["depth"]
    load_frame_to_model(torch_cuda, g, o, model, m2, pose)
    t, R = compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["modelToModel"])
    assert np.max(np.abs(t)) < 2e-3 and np.max(np.abs(R - np.eye(3))) < 2e-3
    cov_o = orc.covariance(g.lastA)
    assert np.allclose(g.getCovariance(), cov_o, rtol=1e-9, atol=1e-12)
    err, cnt = g.getICPResidual(t, R, pose[:3, 3], pose[:3, :3])
    assert cnt > 0 and np.isfinite(err)   # compared with the oracle in test_icp_residual_against_oracle


def test_stopwatch_keys(torch_cuda, orc, synth, pair_small):
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"]))
    g.set_timing(1)
    g.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *FLAG_ROWS["frameToModel"])
    tm = g.timing()
    assert set(tm) >= {"icpStep", "rgbStep", "computeRgbResidual", "so3Step"}  # GPUTest.cpp:280-283 reads these
    assert tm["track"] > 0 and tm["icpStep"] > 0 and tm["rgbStep"] > 0 and tm["so3Step"] > 0


def test_full_track_parity_1080p(torch_cuda, orc, synth):
    """configs[3]: 1920x1080, full pyramid, every iteration against the oracle; the second run on the same handle must
    reproduce the first bit for bit (the reduction does not depend on the launch schedule)."""
    cam = synth.CAM_1080P
    model, live, T = synth.frame_pair(cam, seed=99)
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    runs = []
    for _ in range(2):
        g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
        load_frame_to_model(torch_cuda, g, o, model, live, pose)
        t, R = compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["frameToModel"])
        runs.append((t, R, g.trace()))
    (t0, R0, tr0), (t1, R1, tr1) = runs
    assert np.array_equal(t0, t1) and np.array_equal(R0, R1)
    for a, b in zip(tr0, tr1):
        assert a["icp_count"] == b["icp_count"] and a["rgb_count"] == b["rgb_count"] and a["rgb_sigma"] == b["rgb_sigma"]
        assert np.array_equal(a["icp"], b["icp"]) and np.array_equal(a["rgb"], b["rgb"])
    assert np.max(np.abs(t0 - T[:3, 3])) < 5e-3
