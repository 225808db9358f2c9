"""N1: the processFrame-shaped tracking front end against its CPU restatement (oracle/frontend_ref.py)."""
import numpy as np
import pytest

from conftest import ROOT  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.fixture(scope="module")
def synth():
    from densemonoslam_b200 import synth as s
    return s


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def holed_view(synth, scene, cam, pose, band):
    v = synth.render(scene, cam, pose, seed=5, max_depth=50.0, hole_frac=0.0)
    view = {"vertex": v["vmap4"].copy(), "normal": v["nmap4"].copy(), "image": v["rgba"].copy()}
    if band:  # part of the map not reconstructed yet: empty prediction there (glClear 0)
        y0, y1 = band
        view["vertex"][y0:y1] = 0; view["normal"][y0:y1] = 0; view["image"][y0:y1] = 0
    return view


def test_fill_in_and_dense_fraction_bit_exact(torch_cuda, synth):
    from densemonoslam_b200 import frontend as fe
    from oracle import frontend_ref as ref
    from oracle import oracle as orc
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(3)
    view = holed_view(synth, scene, cam, np.eye(4), (40, 90))
    live = synth.render(scene, cam, synth.pose_from((0.01, 0.0, 0.0), (0.01, 0.0, 0.0)), seed=9, max_depth=50.0)
    filt = orc.depth_bilateral(live["depth"], 3.5)
    c = (cam.cx, cam.cy, cam.fx, cam.fy)
    for pg, pi in ((False, False), (True, False), (False, True)):
        gv, gn, gi = fe.fill_in(dev(torch_cuda, view["vertex"]), dev(torch_cuda, view["normal"]), dev(torch_cuda, view["image"]), dev(torch_cuda, filt),
                                dev(torch_cuda, live["rgba"]), c, pg, pi)
        ov, on, oi = ref.fill_in(view["vertex"], view["normal"], view["image"], filt, live["rgba"], c, pg, pi)
        assert np.array_equal(bits(gv.cpu().numpy()), bits(ov)), "fill vertex"
        gnn, onn = gn.cpu().numpy(), on
        assert np.array_equal(np.isnan(gnn), np.isnan(onn))
        assert np.array_equal(bits(np.nan_to_num(gnn)), bits(np.nan_to_num(onn))), "fill normal"
        assert np.array_equal(gi.cpu().numpy(), oi), "fill image"
    for img in (view["image"], live["rgba"], np.zeros_like(live["rgba"])):
        assert fe.dense_fraction(dev(torch_cuda, img)) == float(ref.dense_fraction(img))
    rng = np.random.RandomState(0)
    for _ in range(20):
        a = synth.pose_from(rng.randn(3) * 0.3, rng.randn(3)); b = a @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.004)
        # rodrigues2 takes acos of a float32 trace: at 4 mrad one float32 ulp of the trace moves the angle by 1.5e-5 rad,
        # i.e. 1.5e-3 of the weight (EF.cpp:955-961); the two restatements agree to that conditioning
        assert abs(fe.velocity_weighting(b, a, 1.5) - ref.velocity_weighting(b.astype(np.float32), a.astype(np.float32), 1.5)) <= 6e-3


def test_process_frame_sequence_matches_restatement(torch_cuda, synth):
    """Six frames through both front ends: first-frame initialisation, dense tracking from the model view, a frame whose
    prediction has a hole (FillIn inputs), the relocalisation gates, and the model-to-model loop-closure check."""
    from densemonoslam_b200.frontend import TrackingFrontEnd
    from oracle.frontend_ref import FrontEndRef
    cam = synth.scaled_camera(synth.CAM_VGA, 2)   # 320 x 240: the loop gate needs > 15000 inliers
    scene = synth.room_scene(11)
    drift = synth.pose_from((0.002, -0.003, 0.001), (0.004, -0.002, 0.003))   # the "old" part of the map is slightly off

    def predictor_np(pose, kind, tick):
        if kind == 2:
            if tick < 4:
                return None   # no inactive part of the map yet
            return holed_view(synth, scene, cam, np.asarray(pose, np.float64) @ drift, None)
        band = (60, 130) if (tick == 3 and kind == 0) else None
        return holed_view(synth, scene, cam, np.asarray(pose, np.float64), band)

    def predictor_gpu(pose, kind, tick):
        v = predictor_np(pose, kind, tick)
        return None if v is None else {k: dev(torch_cuda, a) for k, a in v.items()}

    opts = dict(depthCutoff=3.5, reloc=1, closeLoops=1)
    g = TrackingFrontEnd(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor_gpu, **opts)
    o = FrontEndRef(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor_np, **opts)
    T = np.eye(4)
    rng = np.random.RandomState(4)
    seen_fill, seen_loop = False, False
    for f in range(6):
        T = T @ synth.pose_from((0.002, 0.004 * np.sin(f), 0.001), (0.006, -0.002, 0.004))
        live = synth.render(scene, cam, T, seed=100 + f, max_depth=50.0, depth_noise=0.002)
        prior = (T @ synth.pose_from(rng.randn(3) * 0.003, rng.randn(3) * 0.004)).astype(np.float32)
        rg = g.process(dev(torch_cuda, live["rgba"]), dev(torch_cuda, live["depth"]), prior)
        ro = o.process(live["rgba"], live["depth"], prior)
        for k in ("tick", "tracked", "shouldFillIn", "trackingOk", "lost", "trackingCount", "lastFrameRecovery", "loopChecked", "loopAccepted"):
            assert rg[k] == ro[k], (f, k, rg[k], ro[k])
        assert np.max(np.abs(rg["pose"] - ro["pose"])) <= 1e-5, f      # north_star: pose within 1e-5 m / rad
        assert abs(rg["weighting"] - ro["weighting"]) <= 4e-3, (f, rg["weighting"], ro["weighting"])
        if rg["tracked"]:
            assert rg["denseFraction"] == ro["denseFraction"]
            assert rg["frameToModel"]["lastICPCount"] == ro["frameToModel"].lastICPCount, f
            assert rg["frameToModel"]["lastRGBCount"] == ro["frameToModel"].lastRGBCount, f
        if rg["loopChecked"]:
            assert rg["modelToModel"]["lastICPCount"] == ro["modelToModel"].lastICPCount, f
            assert np.max(np.abs(rg["loopPose"] - ro["loopPose"])) <= 1e-5
            assert np.allclose(rg["loopCovDiag"], ro["loopCovDiag"], rtol=1e-3, atol=1e-12)
        seen_fill |= rg["shouldFillIn"]; seen_loop |= rg["loopChecked"]
    assert seen_fill and seen_loop
