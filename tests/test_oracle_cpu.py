"""CPU-only tests of the oracle (oracle/cpu_odom.cpp): hand-checked quirks of the reference kernels,
self-consistency of the full track, and invariance of the canonical reduction to threading."""
import fractions

import numpy as np

from conftest import small_cam


def f32(x):
    return np.float32(x)


def fma32(a, b, c):
    """Correctly rounded fp32 fma via exact rationals."""
    r = fractions.Fraction(float(a)) * fractions.Fraction(float(b)) + fractions.Fraction(float(c))
    return np.float32(float(r))  # double rounding is harmless here: products of <=8-bit ints by fp32 constants


def test_intensity_matches_reference_expression(orc):
    # cf.cu:651-654: int v = x*0.114f + y*0.299f + z*0.587f  ->  fma(z,.587, fma(x,.114, y*.299)), truncated
    rng = np.random.RandomState(0)
    rgba = rng.randint(0, 256, size=(64, 64, 4)).astype(np.uint8)
    rgba[0, 0] = (255, 255, 255, 255); rgba[0, 1] = (0, 0, 0, 0); rgba[0, 2] = (1, 1, 1, 9)
    out = orc.bgr_to_intensity(rgba)
    for (y, x) in [(0, 0), (0, 1), (0, 2)] + [tuple(rng.randint(0, 64, 2)) for _ in range(300)]:
        r, g, b, _ = [f32(v) for v in rgba[y, x]]
        t = f32(g * f32(0.299))
        v = fma32(b, f32(0.587), fma32(r, f32(0.114), t))
        assert out[y, x] == int(v), (y, x)
    assert out[0, 0] == 255 and out[0, 1] == 0


def test_pyrdown_u8_edge_quirks(orc):
    # exclusive upper bound min(2x+3, cols-1): the last source column/row is never read,
    # and weights are indexed from the window end.
    src = np.full((8, 8), 100, np.uint8)
    src[:, -1] = 255; src[-1, :] = 255
    out = orc.pyrdown_u8_gauss(src)
    assert out.shape == (4, 4) and (out == 100).all()
    # zero taps are skipped, all-zero window -> 0
    z = np.zeros((8, 8), np.uint8)
    assert (orc.pyrdown_u8_gauss(z) == 0).all()
    z[2, 2] = 77
    o = orc.pyrdown_u8_gauss(z)
    assert o[1, 1] == 77 and o[0, 0] == 77 and o[3, 3] == 0


def test_pyrdown_f32_nan_handling(orc):
    src = np.full((8, 8), 2.0, np.float32)
    src[0:4, 0:4] = np.nan
    out = orc.pyrdown_f32_gauss(src)
    assert np.isnan(out[0, 0]) and out[3, 3] == 2.0 and out[1, 1] == 2.0


def test_pyrdown_u16_bilateral(orc):
    src = np.full((8, 8), 1000, np.uint16)
    src[0, 1] = 2000  # differs by >= 90 from the centre -> excluded
    out = orc.pyrdown_u16(src)
    assert (out == 1000).all()
    src[:] = 0
    assert (orc.pyrdown_u16(src) == 0).all()


def test_derivative_border_quirk(orc):
    # interior = true convolution with the flipped 3x3; cf.cu:684-690
    img = np.zeros((5, 6), np.uint8)
    img[:, 3:] = 100
    dx, dy = orc.derivatives(img)
    g = [0.52201, 0.79451, 0.52201]
    assert dx[2, 2] == int(np.float32(100) * np.float32(g[0] + g[1] + g[2]) * 1.0000001) or abs(int(dx[2, 2]) - 183) <= 1
    assert dy[2, 2] == 0
    assert dx[2, 0] == 0


def test_vmap_nmap_and_resize(orc, synth):
    cam = small_cam(synth, 4)
    depth = np.full((cam.height, cam.width), 1500, np.uint16)
    depth[10, 10] = 0
    v = orc.create_vmap(depth, cam.fx, cam.fy, cam.cx, cam.cy, 25.0)
    assert np.isnan(v[0, 10, 10]) and v[2, 5, 5] == np.float32(1.5)
    n = orc.create_nmap(v)
    assert np.isnan(n[0, -1, 0]) and np.isnan(n[0, 0, -1]) and np.isnan(n[0, 10, 9]) and np.isnan(n[0, 9, 10])
    assert abs(abs(n[2, 5, 5]) - 1.0) < 1e-6
    r = orc.resize_map(v, False)
    assert r.shape == (3, cam.height // 2, cam.width // 2) and np.isnan(r[0, 5, 5]) and r[2, 0, 0] == np.float32(1.5)


def _track(orc, cam, model, live, flags, threads=None):
    if threads:
        orc.set_num_threads(threads)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    pose = np.eye(4, dtype=np.float32)
    o.initFirstRGB(model["rgba"])
    o.initICPModel(model["vmap4"], model["nmap4"], pose); o.initRGBModel(model["rgba"])
    o.initICP(live["depth"], 25.0); o.initRGB(live["rgba"])
    t, R, res = o.getIncrementalTransformation(np.zeros(3), np.eye(3), *flags)
    return t, R, res, o.log()


def test_track_recovers_rendered_motion(orc, synth, pair_small):
    cam, model, live, T = pair_small
    t, R, res, log = _track(orc, cam, model, live, (False, 10.0, True, False, True, False))
    assert list(res.iters) == [10, 5, 4]
    assert np.max(np.abs(t - T[:3, 3])) < 6e-3, (t, T[:3, 3])
    assert np.max(np.abs(R - T[:3, :3])) < 3e-3
    assert res.lastICPCount > 0.5 * cam.width * cam.height
    # identity motion -> (near) zero update
    t0, R0, _, _ = _track(orc, cam, model, {"depth": model["depth"], "rgba": model["rgba"]}, (False, 10.0, True, False, False, False))
    assert np.max(np.abs(t0)) < 2e-3 and np.max(np.abs(R0 - np.eye(3))) < 2e-3


def test_reduction_is_schedule_independent(orc, synth, pair_small):
    """The exact fixed-point accumulation makes the sums independent of the thread count."""
    cam, model, live, T = pair_small
    flags = (False, 10.0, True, False, True, False)
    n = orc.num_threads()
    a = _track(orc, cam, model, live, flags, threads=1)
    b = _track(orc, cam, model, live, flags, threads=max(2, n))
    orc.set_num_threads(n)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for x, y in zip(a[3], b[3]):
        assert np.array_equal(x["icp"], y["icp"]) and np.array_equal(x["rgb"], y["rgb"]) and x["icp_count"] == y["icp_count"]


def test_flag_schedule(orc, synth, pair_small):
    cam, model, live, T = pair_small
    # RO.cpp:387-389
    assert list(_track(orc, cam, model, live, (False, 10.0, False, True, False, False))[2].iters) == [3, 0, 0]
    assert list(_track(orc, cam, model, live, (False, 10.0, False, False, False, False))[2].iters) == [10, 0, 0]
    r = _track(orc, cam, model, live, (False, 100.0, True, False, False, True))[2]
    assert list(r.iters) == [50, 50, 50] and r.lastRGBCount == 0
    r = _track(orc, cam, model, live, (True, 10.0, True, False, True, False))[2]
    assert r.lastICPCount == 0 and r.lastRGBCount > 0 and all(i <= n for i, n in zip(r.iters, (10, 5, 4)))


def test_guard_rejects_large_translation(orc, synth):
    # RO.cpp:589-593: |tcurr - tprev| > 0.3 m with an RGB term -> input pose returned
    cam = small_cam(synth, 4)
    model, live, T = synth.frame_pair(cam, seed=3, tvec=(0.9, 0.0, 0.0), rvec=(0, 0, 0))
    t, R, res, _ = _track(orc, cam, model, live, (False, 10.0, True, False, False, False))
    assert np.linalg.norm(t) <= 0.3 + 1e-6


def test_depth_prefilter_oracle(orc):
    # depth_bilateral.frag:33-36: < 300 or > maxD*1000 -> 0; a constant patch is a fixed point; exp_neg accuracy
    import math
    d = np.full((40, 50), 1500, np.uint16); d[5, 5] = 100; d[10, 10] = 0; d[20, 20] = 5000
    f = orc.depth_bilateral(d, 3.0)
    assert f[5, 5] == 0 and f[10, 10] == 0 and f[20, 20] == 0 and f[0, 0] == 1500 and f[30, 30] == 1500
    m = orc.depth_metric(d, 3.0)
    assert m[0, 0] == np.float32(1.5) and m[5, 5] == 0 and m[20, 20] == 0
    for t in (0.0, 0.3, 1.0, 7.7, 40.0):
        assert abs(orc.exp_neg(t) - math.exp(-t)) <= 1e-5 * math.exp(-t)
    rng = np.random.RandomState(0)
    noisy = (1500 + rng.randint(-20, 21, size=(40, 50))).astype(np.uint16)
    g = orc.depth_bilateral(noisy, 3.0).astype(int)
    assert np.abs(g - 1500).mean() < np.abs(noisy.astype(int) - 1500).mean()   # it smooths


def test_frontend_restatement_pieces():
    """oracle/frontend_ref.py (N1): FillIn selection rules, denseEnough sampling, rodrigues2, velocity weighting."""
    from oracle import frontend_ref as ref
    H, W = 40, 60
    mv = np.zeros((H, W, 4), np.float32); mn = np.zeros((H, W, 4), np.float32); mi = np.zeros((H, W, 4), np.uint8)
    mv[:20, :, :3] = (0.1, 0.2, 1.5); mn[:20, :, :3] = (0.0, 0.0, -1.0); mi[:20] = (10, 20, 30, 255)
    depth = np.full((H, W), 2000, np.uint16); live = np.full((H, W, 4), 77, np.uint8)
    cam = (29.5, 19.5, 50.0, 50.0)
    ov, on, oi = ref.fill_in(mv, mn, mi, depth, live, cam)
    assert np.array_equal(ov[:20], mv[:20]) and np.array_equal(on[:20], mn[:20]) and np.array_equal(oi[:20], mi[:20])
    assert np.all(ov[20:, :, 2] == np.float32(2.0)) and np.all(oi[20:] == 77)
    assert ov[30, 40, 0] == (np.float32(40) - np.float32(29.5)) * np.float32(2.0) * (np.float32(1) / np.float32(50))
    assert np.allclose(on[25, 10, :3], (0, 0, 1), atol=1e-6)   # fronto-parallel plane: cross(dx, dy) = +z
    ov2, _, oi2 = ref.fill_in(mv, mn, mi, depth, live, cam, pass_geom=True, pass_img=True)
    assert np.all(ov2[..., 2] == np.float32(2.0)) and np.all(oi2 == 77)
    assert ref.dense_fraction(mi) <= np.float32(0.5) and ref.dense_fraction(live) == np.float32(1.0)
    rv = np.array([0.01, -0.02, 0.015])
    from densemonoslam_b200 import synth
    T = synth.pose_from(rv, (0, 0, 0))
    assert np.allclose(ref.rodrigues2(T[:3, :3]), rv, atol=1e-6)
    a = synth.pose_from((0.1, 0.2, 0.3), (1, 2, 3)).astype(np.float32)
    assert ref.velocity_weighting(a, a, 2.0) == 2.0                      # no motion: full weight
    b = (a @ synth.pose_from((0, 0, 0), (0.005, 0, 0))).astype(np.float32)
    assert abs(ref.velocity_weighting(b, a) - 0.5) < 1e-3               # half of `largest`
    c = (a @ synth.pose_from((0.2, 0, 0), (0, 0, 0))).astype(np.float32)
    assert ref.velocity_weighting(c, a) == 0.5                           # clamped at minWeight


def test_nid_restatement_pieces():
    """oracle/nid_ref.py (N4): NaN-comparison quirk of the key-frame pick, bin rules, entropy identities."""
    from oracle import nid_ref as nr
    nan = np.float32(np.nan)
    dk = np.array([[1.0, nan, 2.0, nan]], np.float32); dko = np.array([[2.0, 1.0, nan, nan]], np.float32)
    assert nr.pick_active(dk, dko).tolist() == [[True, False, False, False]]
    ik = np.array([[10, 20, 30, 40]], np.uint8); iko = np.array([[200, 210, 220, 230]], np.uint8); ic = np.array([[0, 255, 128, 64]], np.uint8)
    h = nr.hist_img(ik, iko, dk, dko, ic, 64)
    assert h.sum() == 4 and h[0, 10 // 4] == 1 and h[255 // 4, 210 // 4] == 1 and h[128 // 4, 220 // 4] == 1
    hd = nr.hist_depth(dk, dko, np.array([[1.0, nan, 2.9, 5.0]], np.float32), 500, 3000.0)
    assert hd[int(1000 / 6), int(1000 / 6)] == 1      # 1 m vs 1 m
    assert hd[0, int(1000 / 6)] == 1                  # NaN live depth -> bin 0 (device float->int of NaN)
    assert hd.sum() == 3                              # 5 m live depth falls outside the histogram: dropped
    rng = np.random.RandomState(0)
    a = rng.randint(0, 256, (60, 80)).astype(np.uint8)
    z = np.ones((60, 80), np.float32)
    same = nr.nid_from_hist(nr.hist_img(a, a, z, z, a, 64), a.size)
    indep = nr.nid_from_hist(nr.hist_img(a, a, z, z, rng.randint(0, 256, (60, 80)).astype(np.uint8), 64), a.size)
    assert abs(float(same)) < 1e-5 and 0.7 < float(indep) <= 1.0   # identical images: distance 0; independent: near 1


def test_splat_restatement_pieces():
    """oracle/splat_ref.py (N3): disk footprint, depth order / draw-order ties, visibility rules of splat.vert:46-50."""
    from oracle import splat_ref as sr
    cam = (31.5, 23.5, 60.0, 60.0); rows, cols = 48, 64
    def surfel(x, y, z, rad, col, conf=10.0, t=5.0, init=5.0, n=(0, 0, -1)):
        s = np.zeros(15, np.float32); s[0:4] = (x, y, z, conf); s[4] = col; s[6] = init; s[8:11] = (t, -3, -3); s[11:14] = n; s[14] = rad
        return s
    I = np.eye(4, dtype=np.float32)
    args = dict(t_inv=I, cam=cam, rows=rows, cols=cols, maxDepth=5.0, confThreshold=5.0, time=6, timeIdx=0, maxTime=6, timeDelta=200, active=1)
    # one fronto-parallel surfel at 2 m with radius 0.1 m: a disk of radius 0.1 * 60 / 2 = 3 px around (31.5, 23.5)
    img, ver, nor, tim = sr.predict(np.stack([surfel(0, 0, 2.0, 0.1, (200 << 16) + (100 << 8) + 50)]), **args)
    yy, xx = np.nonzero(ver[..., 2] > 0)
    r = np.hypot(xx + 0.5 - 31.5, yy + 0.5 - 23.5)
    assert r.max() <= 3.0 + 1e-3 and len(r) == int((np.hypot(*np.meshgrid(np.arange(64) + 0.5 - 31.5, np.arange(48) + 0.5 - 23.5)) <= 3.0).sum())
    assert img[23, 31].tolist() == [200, 100, 50, 255] and tim[23, 31] == 5 and ver[23, 31, 2] == np.float32(2.0) and ver[23, 31, 3] == 10.0
    # nearer surfel wins; equal depth: the one drawn first wins
    near = surfel(0, 0, 1.5, 0.02, 7); same_a = surfel(0.2, 0, 2.0, 0.05, 11); same_b = surfel(0.2, 0, 2.0, 0.05, 12)
    img, ver, _, _ = sr.predict(np.stack([surfel(0, 0, 2.0, 0.1, 9), near, same_a, same_b]), **args)
    assert img[23, 31, 2] == 7 and ver[23, 31, 2] == np.float32(1.5)
    px = int(31.5 + 0.2 * 60 / 2.0)
    assert img[23, px, 2] == 11
    # visibility: low confidence, beyond maxDepth, stale (time - t > timeDelta), newer than maxTime, unseen (-3) in ACTIVE mode
    for bad in (surfel(0, 0, 2, 0.1, 1, conf=1.0), surfel(0, 0, 6, 0.1, 1), surfel(0, 0, 2, 0.1, 1, t=7.0)):
        assert not (sr.predict(np.stack([bad]), **args)[1][..., 2] > 0).any()
    assert not (sr.predict(np.stack([surfel(0, 0, 2, 0.1, 1, t=1.0)]), **{**args, "timeDelta": 3})[1][..., 2] > 0).any()
    unseen = surfel(0, 0, 2, 0.1, 1, t=-3.0)
    assert not (sr.predict(np.stack([unseen]), **args)[1][..., 2] > 0).any()
    assert (sr.predict(np.stack([unseen]), **{**args, "active": 0})[1][..., 2] > 0).any()   # INACTIVE keeps -3 surfels (splat.vert:46)


def test_synthetic_surfel_map_matches_the_checker_helper(synth):
    """bench.py's front-end leg builds its surfel map with synth.surfels_from_view (the product side may not touch oracle/);
    it must lay the 15 floats out exactly as the helper the splat parity tests use."""
    from oracle import splat_ref as sr
    cam = synth.scaled_camera(synth.CAM_VGA, 4)
    T = synth.pose_from((0.02, -0.03, 0.01), (0.1, -0.05, 0.2))
    view = synth.render(synth.room_scene(5), cam, T, seed=9, hole_frac=0.02)
    a = synth.surfels_from_view(view, T, cam, time=3, conf=10.0)
    b = sr.surfels_from_view(view["vmap4"], view["nmap4"], view["rgba"], T, (cam.cx, cam.cy, cam.fx, cam.fy), time=3, sensor=0, conf=10.0)
    assert a.shape == b.shape and a.shape[1] == synth.SURFEL_FLOATS == sr.SURFEL_FLOATS and len(a) > 1000
    assert np.array_equal(a, b)
