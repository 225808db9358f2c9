"""N3: surfel-splat model-view synthesis against its CPU restatement (oracle/splat_ref.py)."""
import numpy as np
import pytest

from conftest import ROOT  # noqa: F401

pytestmark = pytest.mark.gpu


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def make_map(synth, sr, cam, scene):
    """Two overlapping views fused naively into one surfel list (second view drawn later = loses depth ties); a band of
    the first view is 'old' (time 1), the rest recent (time 40); some surfels unstable (low confidence) or never seen by
    this sensor (-3)."""
    c = (cam.cx, cam.cy, cam.fx, cam.fy)
    va = synth.render(scene, cam, np.eye(4), seed=1, max_depth=50.0, hole_frac=0.0)
    pb = synth.pose_from((0.0, 0.05, 0.0), (0.15, 0.0, 0.0))
    vb = synth.render(scene, cam, pb, seed=2, max_depth=50.0, hole_frac=0.0)
    A = sr.surfels_from_view(va["vmap4"], va["nmap4"], va["rgba"], np.eye(4), c, time=40)
    B = sr.surfels_from_view(vb["vmap4"], vb["nmap4"], vb["rgba"], pb, c, time=40)
    S = np.concatenate([A, B])
    rng = np.random.RandomState(0)
    old = rng.rand(len(S)) < 0.3
    S[old, 6] = 1; S[old, 8] = 1
    S[rng.rand(len(S)) < 0.1, 3] = 2.0      # below the confidence threshold
    S[rng.rand(len(S)) < 0.05, 8] = -3.0    # not observed by sensor 0
    return S, c


@pytest.mark.parametrize("active,time,maxTime,timeDelta", [(1, 41, 41, 200), (1, 41, 41, 20), (0, 0, 21, 200)])
def test_splat_predict_bit_exact(active, time, maxTime, timeDelta):
    import torch
    from densemonoslam_b200 import splat, synth
    from oracle import splat_ref as sr
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    S, c = make_map(synth, sr, cam, synth.room_scene(5))
    pose = synth.pose_from((0.01, -0.03, 0.004), (0.06, 0.01, -0.02)).astype(np.float32)
    t_inv = splat.invert_pose(pose)
    assert np.max(np.abs(t_inv @ pose - np.eye(4))) < 1e-6
    gi, gv, gn, gt = splat.combinedPredict(dev(torch, S), pose, c, cam.height, cam.width, 5.0, 5.0, time, 0, maxTime, timeDelta, active, t_inv=t_inv)
    oi, ov, on, ot = sr.predict(S, t_inv, c, cam.height, cam.width, 5.0, 5.0, time, 0, maxTime, timeDelta, active)
    assert np.array_equal(gt.cpu().numpy().astype(np.uint32), ot)
    assert np.array_equal(gi.cpu().numpy(), oi)
    assert np.array_equal(bits(gv.cpu().numpy()), bits(ov))
    assert np.array_equal(bits(gn.cpu().numpy()), bits(on))
    cover = float((ov[..., 2] > 0).mean())
    assert cover > 0.2, cover


def test_front_end_tracks_against_splatted_map():
    """Closes the loop without GL: the front end's model views come from dms_splat_predict of a surfel map."""
    import torch
    from densemonoslam_b200 import splat, synth
    from densemonoslam_b200.frontend import TrackingFrontEnd
    from oracle import splat_ref as sr
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(8)
    c = (cam.cx, cam.cy, cam.fx, cam.fy)
    v0 = synth.render(scene, cam, np.eye(4), seed=1, max_depth=50.0, hole_frac=0.0)
    S = dev(torch, sr.surfels_from_view(v0["vmap4"], v0["nmap4"], v0["rgba"], np.eye(4), c, time=1))

    def predictor(pose, kind, tick):
        if kind == 2:
            return None
        img, ver, nor, _ = splat.combinedPredict(S, pose, c, cam.height, cam.width, 3.5, 5.0, tick, 0, tick, 200, splat.ACTIVE)
        return {"vertex": ver, "normal": nor, "image": img}

    fe = TrackingFrontEnd(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor, depthCutoff=3.5, closeLoops=0)
    # combo_splat.frag evaluates the model at gl_FragCoord = pixel + 0.5 while the live maps use integer pixels
    # (cf.cu:106-128): a constant half-pixel offset between model and live, hence a constant pose offset.  The
    # frame-to-frame motion is free of it, so that is what is compared with the ground truth.
    T = np.eye(4)
    rng = np.random.RandomState(3)
    gt, est, pri = [], [], []
    for f in range(5):
        if f:
            T = T @ synth.pose_from((0.003, -0.004, 0.001), (0.008, -0.003, 0.004))
        live = synth.render(scene, cam, T, seed=50 + f, max_depth=50.0, depth_noise=0.001)
        prior = (T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)).astype(np.float32) if f else np.eye(4, dtype=np.float32)
        r = fe.process(dev(torch, live["rgba"]), dev(torch, live["depth"]), prior)
        if f:
            assert r["tracked"] and r["frameToModel"]["lastICPCount"] > 20000
            gt.append(T.copy()); est.append(r["pose"].astype(np.float64)); pri.append(prior.astype(np.float64))
    rel = lambda a, b: np.linalg.inv(a) @ b
    e_est = [np.linalg.norm(rel(est[i], est[i + 1])[:3, 3] - rel(gt[i], gt[i + 1])[:3, 3]) for i in range(len(gt) - 1)]
    e_pri = [np.linalg.norm(rel(pri[i], pri[i + 1])[:3, 3] - rel(gt[i], gt[i + 1])[:3, 3]) for i in range(len(gt) - 1)]
    assert np.mean(e_est) < 0.3 * np.mean(e_pri), (e_est, e_pri)


def test_splat_edge_cases():
    """Empty map, a surfel right in front of the camera (sprite larger than the image), surfels on the image border,
    behind the camera and edge-on; ragged image size."""
    import torch
    from densemonoslam_b200 import splat
    from oracle import splat_ref as sr
    rows, cols = 37, 53
    c = (25.7, 18.2, 40.0, 41.0)
    I = np.eye(4, dtype=np.float32)
    kw = dict(depthCutoff=5.0, confThreshold=1.0, time=3, timeIdx=1, maxTime=3, timeDelta=10)
    empty = torch.zeros((0, 15), dtype=torch.float32, device="cuda")
    gi, gv, gn, gt = splat.combinedPredict(empty, I, c, rows, cols, predictionType=1, t_inv=I, **kw)
    assert not gi.any() and not gv.any() and not gn.any() and not gt.any()

    def surfel(x, y, z, rad, n, col=0x102030):
        s = np.zeros(15, np.float32); s[0:4] = (x, y, z, 9.0); s[4] = col; s[6] = 2; s[8:11] = (-3, 2, -3); s[11:14] = n; s[14] = rad
        return s
    S = np.stack([surfel(0.0, 0.0, 0.05, 0.5, (0, 0, -1), 0x0000ff),            # fills the whole view
                  surfel(-1.2, 0.3, 2.0, 0.2, (0.3, 0, -0.95)),                  # centre just outside the left border: clipped
                  surfel(1.25, -0.8, 2.0, 0.15, (0, 0.2, -0.98), 0x00ff00),      # near the top-right corner
                  surfel(0.1, 0.1, -1.0, 0.1, (0, 0, -1)),                       # behind the camera
                  surfel(0.3, 0.2, 0.04, 0.05, (1, 0, 0), 0xff0000),             # edge-on, nearer than the big one
                  surfel(-0.2, 0.1, 0.03, 0.004, (0.1, -0.2, -0.97), 0xffff00)])  # small, nearest
    gi, gv, gn, gt = splat.combinedPredict(torch.from_numpy(S).cuda(), I, c, rows, cols, predictionType=1, t_inv=I, **kw)
    oi, ov, on, ot = sr.predict(S, I, c, rows, cols, 5.0, 1.0, 3, 1, 3, 10, 1)
    assert np.array_equal(gi.cpu().numpy(), oi) and np.array_equal(gt.cpu().numpy().astype(np.uint32), ot)
    assert np.array_equal(bits(gv.cpu().numpy()), bits(ov)) and np.array_equal(bits(gn.cpu().numpy()), bits(on))
    assert (ov[..., 2] > 0).mean() > 0.9   # the big surfel covers the image
